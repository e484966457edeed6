/* Drop-in shim: the reference's include/hittable.h surface is declared in rtc_api.h (see its header comment). */
#ifndef RTC_SHIM_HITTABLE_H
#define RTC_SHIM_HITTABLE_H
#include "rtc_api.h"
#endif

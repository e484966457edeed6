/* Drop-in shim: the reference's include/utils.h surface is declared in rtc_api.h (see its header comment). */
#ifndef RTC_SHIM_UTILS_H
#define RTC_SHIM_UTILS_H
#include "rtc_api.h"
#endif

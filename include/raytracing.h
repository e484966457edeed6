/* Drop-in shim: the reference's include/raytracing.h surface is declared in rtc_api.h (see its header comment). */
#ifndef RTC_SHIM_RAYTRACING_H
#define RTC_SHIM_RAYTRACING_H
#include "rtc_api.h"
#endif

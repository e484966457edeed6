/* Drop-in shim: the reference's include/pcg32.h surface is declared in rtc_api.h (see its header comment). */
#ifndef RTC_SHIM_PCG32_H
#define RTC_SHIM_PCG32_H
#include "rtc_api.h"
#endif

/* Drop-in shim: the reference's include/tiff.h surface is declared in rtc_api.h (see its header comment). */
#ifndef RTC_SHIM_TIFF_H
#define RTC_SHIM_TIFF_H
#include "rtc_api.h"
#endif

/* Drop-in shim: the reference's include/texture.h surface is declared in rtc_api.h (see its header comment). */
#ifndef RTC_SHIM_TEXTURE_H
#define RTC_SHIM_TEXTURE_H
#include "rtc_api.h"
#endif

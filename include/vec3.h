/* Drop-in shim: the reference's include/vec3.h surface is declared in rtc_api.h (see its header comment). */
#ifndef RTC_SHIM_VEC3_H
#define RTC_SHIM_VEC3_H
#include "rtc_api.h"
#endif

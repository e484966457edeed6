/* Drop-in shim: the reference's include/material.h surface is declared in rtc_api.h (see its header comment). */
#ifndef RTC_SHIM_MATERIAL_H
#define RTC_SHIM_MATERIAL_H
#include "rtc_api.h"
#endif

/*
 * tiff_out.c -- baseline TIFF writer behind write_tiff() (rtc_api.h).
 *
 * Produces the same bytes as the reference's writer (/root/reference/src/tiff.c:15-49): little-endian header,
 * one IFD of 11 entries, BitsPerSample / XResolution / YResolution stored right after the IFD, then a single
 * uncompressed strip.  Built here from a table of entries instead of a sequence of calls.
 */
#include "rtc_api.h"
#include <string.h>

enum { TIFF_SHORT = 3, TIFF_LONG = 4, TIFF_RATIONAL = 5 };

static void put_u16(uint8_t *p, uint32_t v) {
  p[0] = (uint8_t)(v & 0xff);
  p[1] = (uint8_t)((v >> 8) & 0xff);
}
static void put_u32(uint8_t *p, uint32_t v) {
  put_u16(p, v & 0xffff);
  put_u16(p + 2, v >> 16);
}

int write_tiff(FILE *f, int width, int height, int n_channels, uint8_t *buffer) {
  if (n_channels != 1 && n_channels != 3)
    return 1;

  enum { N_ENTRIES = 11, IFD_AT = 8 };
  const uint32_t extra_at = IFD_AT + 2 + N_ENTRIES * 12 + 4; /* first byte after the IFD */
  const uint32_t xres_at = extra_at + 2u * (uint32_t)n_channels;
  const uint32_t yres_at = xres_at + 8;
  const uint32_t strip_at = yres_at + 8;
  const uint32_t n_bytes = (uint32_t)width * (uint32_t)height * (uint32_t)n_channels;

  const struct {
    uint16_t tag, type;
    uint32_t count, value;
  } entry[N_ENTRIES] = {
      {0x100, TIFF_SHORT, 1, (uint32_t)width},                  /* ImageWidth */
      {0x101, TIFF_SHORT, 1, (uint32_t)height},                 /* ImageLength */
      {0x103, TIFF_SHORT, 1, 1},                                /* Compression: none */
      {0x102, TIFF_SHORT, n_channels == 1 ? 1u : 3u, n_channels == 1 ? 8u : extra_at}, /* BitsPerSample */
      {0x106, TIFF_SHORT, 1, n_channels == 1 ? 1u : 2u},        /* Photometric: min-is-black / RGB */
      {0x111, TIFF_SHORT, 1, strip_at},                         /* StripOffsets */
      {0x115, TIFF_SHORT, 1, (uint32_t)n_channels},             /* SamplesPerPixel */
      {0x116, TIFF_SHORT, 1, (uint32_t)height},                 /* RowsPerStrip: one strip */
      {0x117, TIFF_LONG, 1, n_bytes},                           /* StripByteCounts */
      {0x11A, TIFF_RATIONAL, 1, xres_at},                       /* XResolution */
      {0x11B, TIFF_RATIONAL, 1, yres_at},                       /* YResolution */
  };

  uint8_t head[IFD_AT + 2 + N_ENTRIES * 12 + 4 + 6 + 16];
  memset(head, 0, sizeof(head));
  head[0] = 'I', head[1] = 'I';
  put_u16(head + 2, 42);
  put_u32(head + 4, IFD_AT);
  put_u16(head + IFD_AT, N_ENTRIES);
  for (int i = 0; i < N_ENTRIES; i++) {
    uint8_t *e = head + IFD_AT + 2 + 12 * i;
    put_u16(e, entry[i].tag);
    put_u16(e + 2, entry[i].type);
    put_u32(e + 4, entry[i].count);
    put_u32(e + 8, entry[i].value); /* the reference stores every value as a 32-bit little-endian word */
  }
  /* next-IFD offset stays 0 */
  uint8_t *extra = head + extra_at;
  for (int c = 0; c < n_channels; c++)
    put_u16(extra + 2 * c, 8);
  put_u32(head + xres_at, 72), put_u32(head + xres_at + 4, 1);
  put_u32(head + yres_at, 72), put_u32(head + yres_at + 4, 1);

  if (fwrite(head, 1, strip_at, f) != strip_at)
    return 1;
  if (fwrite(buffer, 1, n_bytes, f) != n_bytes)
    return 1;
  return 0;
}

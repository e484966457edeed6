/*
 * image_io.c -- image-texture ingest that does not depend on the object model: a binary-PPM reader and the
 * procedural stand-in for the reference's earthmap.jpg (the asset is not in the reference repository; its
 * scenes 3 and 7 load it from the working directory, /root/reference/src/main.c:104,243).
 * Kept in its own file so that oracle/ref_tool.c can link it next to the reference's own constructors.
 */
#include "rtc_host.h"

/* The procedural stand-in for earthmap.jpg: 1024x512, R ramps with x, G with y, B is a 32-pixel checker.
 * oracle/ writes the same texels as a binary PPM for the reference build, so CPU and GPU see identical data. */
void rtc_synthetic_image(int width, int height, uint8_t *rgb) {
  for (int y = 0; y < height; y++)
    for (int x = 0; x < width; x++) {
      uint8_t *px = rgb + 3 * ((size_t)y * width + x);
      px[0] = (uint8_t)((x * 255) / (width - 1));
      px[1] = (uint8_t)((y * 255) / (height - 1));
      px[2] = (((x >> 5) + (y >> 5)) & 1) ? 230 : 40;
    }
}

static int ppm_next_int(FILE *f, int *out) {
  int c = fgetc(f);
  for (;;) { /* skip whitespace and # comments */
    while (c == ' ' || c == '\t' || c == '\n' || c == '\r')
      c = fgetc(f);
    if (c != '#')
      break;
    while (c != '\n' && c != EOF)
      c = fgetc(f);
  }
  if (c < '0' || c > '9')
    return -1;
  int v = 0;
  while (c >= '0' && c <= '9') {
    v = v * 10 + (c - '0');
    c = fgetc(f);
  }
  *out = v; /* the single whitespace byte after the number has been consumed */
  return 0;
}

/* Binary PPM (P6, maxval 255).  Returns a malloc'ed RGB8 buffer or NULL. */
uint8_t *rtc_load_ppm(const char *filename, int *width, int *height) {
  FILE *f = fopen(filename, "rb");
  if (f == NULL)
    return NULL;
  uint8_t *pixels = NULL;
  int w = 0, h = 0, maxval = 0;
  if (fgetc(f) == 'P' && fgetc(f) == '6' && ppm_next_int(f, &w) == 0 && ppm_next_int(f, &h) == 0 &&
      ppm_next_int(f, &maxval) == 0 && maxval == 255 && w > 0 && h > 0) {
    size_t n = (size_t)w * h * 3;
    pixels = my_malloc(n);
    if (fread(pixels, 1, n, f) != n) {
      free(pixels);
      pixels = NULL;
    }
  }
  fclose(f);
  if (pixels != NULL) {
    *width = w;
    *height = h;
  }
  return pixels;
}


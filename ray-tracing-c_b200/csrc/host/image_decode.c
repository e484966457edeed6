/*
 * image_decode.c -- decoders for the image texture's on-disk formats: PNG and baseline JPEG, to 8-bit RGB.
 *
 * The reference loads its image texture with `stbi_load(filename, &w, &h, NULL, 3)` (/root/reference/src/texture.c:38-42;
 * stb_image.h v2.29, a third-party single-header library it vendors as a git submodule) and asserts on failure.  The
 * texels feed Image_value (texture.c:28-37), so a drop-in must produce THE SAME BYTES from the same file:
 *   PNG   lossless -- any correct decoder agrees; what has to match is stb's conversion to 3 channels (alpha dropped, not
 *         composited; gray replicated; 16-bit samples truncated to their high byte; 1/2/4-bit gray scaled by 255/85/17);
 *   JPEG  lossy -- the bytes depend on the decoder's arithmetic, so this file restates the published algorithms stb
 *         uses: the IJG "islow" integer inverse DCT at 12-bit constants with its two rounding shifts (>>10, >>17 with
 *         the +128 level shift folded in), JFIF-centred "fancy" chroma upsampling (3:1 taps, +2>>2 / +8>>4 rounding) and
 *         the 20-bit fixed-point YCbCr->RGB conversion.  Written from scratch around those formulas (own bit reader,
 *         own Huffman tables, whole-plane upsampling instead of stb's row callbacks).
 * tests/test_image_decode.py compares both decoders with stb's output byte for byte (fixtures made by
 * `oracle/_ref/ref_tool image`, i.e. the reference's own Image_init).
 *
 * Not decoded (a loud failure, never a silent substitute): progressive and arithmetic-coded JPEG, 12-bit JPEG, CMYK/YCCK
 * JPEG, PNG with a malformed stream.
 */
#include "rtc_host.h"
#include <string.h>

static const char *g_why = "";
const char *rtc_image_error(void) { return g_why; }
static void *fail_null(const char *why) {
  g_why = why;
  return NULL;
}

static uint32_t be32(const uint8_t *p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }
static uint32_t be16(const uint8_t *p) { return ((uint32_t)p[0] << 8) | p[1]; }

/* ================================================================== inflate (RFC 1950 / 1951) ==================== */

typedef struct {
  const uint8_t *in;
  size_t n_in, at;
  uint32_t bits;
  int n_bits;
  uint8_t *out;
  size_t n_out, cap;
  int bad;
} Inflate;

typedef struct { /* canonical Huffman code: how many codes of each length, and the symbols ordered by code */
  uint16_t count[16];
  uint16_t symbol[288];
} Huff;

static uint32_t take_bits(Inflate *s, int n) {
  while (s->n_bits < n) {
    if (s->at >= s->n_in) {
      s->bad = 1;
      return 0;
    }
    s->bits |= (uint32_t)s->in[s->at++] << s->n_bits;
    s->n_bits += 8;
  }
  uint32_t v = s->bits & ((n < 32 ? (1u << n) : 0u) - 1u);
  s->bits >>= n;
  s->n_bits -= n;
  return v;
}
static void huff_build(Huff *h, const uint8_t *lengths, int n) {
  memset(h->count, 0, sizeof(h->count));
  for (int i = 0; i < n; i++)
    h->count[lengths[i]]++;
  h->count[0] = 0;
  uint16_t first[16];
  first[1] = 0;
  for (int len = 1; len < 15; len++)
    first[len + 1] = (uint16_t)(first[len] + h->count[len]);
  for (int i = 0; i < n; i++)
    if (lengths[i])
      h->symbol[first[lengths[i]]++] = (uint16_t)i;
}
static int huff_decode(Inflate *s, const Huff *h) {
  int code = 0, first = 0, index = 0;
  for (int len = 1; len <= 15; len++) {
    code |= (int)take_bits(s, 1);
    int count = h->count[len];
    if (code - count < first)
      return h->symbol[index + (code - first)];
    index += count;
    first = (first + count) << 1;
    code <<= 1;
  }
  s->bad = 1;
  return -1;
}
static void put_byte(Inflate *s, uint8_t b) {
  if (s->n_out == s->cap) {
    size_t cap = s->cap ? s->cap * 2 : 1 << 16;
    uint8_t *grown = realloc(s->out, cap);
    if (grown == NULL) {
      s->bad = 1;
      return;
    }
    s->out = grown, s->cap = cap;
  }
  s->out[s->n_out++] = b;
}
static void inflate_codes(Inflate *s, const Huff *lit, const Huff *dist) {
  static const uint16_t len_base[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
  static const uint8_t len_extra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
  static const uint16_t dist_base[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
  static const uint8_t dist_extra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};
  while (!s->bad) {
    int sym = huff_decode(s, lit);
    if (sym < 0)
      return;
    if (sym < 256) {
      put_byte(s, (uint8_t)sym);
    } else if (sym == 256) {
      return;
    } else {
      sym -= 257;
      if (sym >= 29) {
        s->bad = 1;
        return;
      }
      size_t len = len_base[sym] + take_bits(s, len_extra[sym]);
      int d = huff_decode(s, dist);
      if (d < 0 || d >= 30) {
        s->bad = 1;
        return;
      }
      size_t back = dist_base[d] + take_bits(s, dist_extra[d]);
      if (back > s->n_out) {
        s->bad = 1;
        return;
      }
      for (size_t k = 0; k < len && !s->bad; k++)
        put_byte(s, s->out[s->n_out - back]);
    }
  }
}
/* zlib stream -> malloc'ed bytes */
static uint8_t *zlib_inflate(const uint8_t *in, size_t n_in, size_t *n_out) {
  if (n_in < 2 || (in[0] & 15) != 8 || (((unsigned)in[0] << 8) | in[1]) % 31 != 0 || (in[1] & 32))
    return fail_null("PNG: bad zlib header");
  Inflate s = {in, n_in, 2, 0, 0, NULL, 0, 0, 0};
  int last;
  do {
    last = (int)take_bits(&s, 1);
    uint32_t type = take_bits(&s, 2);
    if (type == 0) { /* stored */
      s.bits = 0, s.n_bits = 0;
      if (s.at + 4 > s.n_in) {
        s.bad = 1;
        break;
      }
      size_t len = s.in[s.at] | ((size_t)s.in[s.at + 1] << 8);
      s.at += 4;
      if (s.at + len > s.n_in) {
        s.bad = 1;
        break;
      }
      for (size_t k = 0; k < len; k++)
        put_byte(&s, s.in[s.at++]);
    } else if (type == 1) { /* fixed codes */
      uint8_t lengths[288 + 30];
      for (int i = 0; i < 288; i++)
        lengths[i] = i < 144 ? 8 : i < 256 ? 9 : i < 280 ? 7 : 8;
      Huff lit, dist;
      huff_build(&lit, lengths, 288);
      for (int i = 0; i < 30; i++)
        lengths[i] = 5;
      huff_build(&dist, lengths, 30);
      inflate_codes(&s, &lit, &dist);
    } else if (type == 2) { /* dynamic codes */
      static const uint8_t order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
      int n_lit = (int)take_bits(&s, 5) + 257, n_dist = (int)take_bits(&s, 5) + 1, n_code = (int)take_bits(&s, 4) + 4;
      uint8_t lengths[288 + 32];
      memset(lengths, 0, sizeof(lengths));
      for (int i = 0; i < n_code; i++)
        lengths[order[i]] = (uint8_t)take_bits(&s, 3);
      Huff code;
      huff_build(&code, lengths, 19);
      memset(lengths, 0, sizeof(lengths));
      int i = 0;
      while (i < n_lit + n_dist && !s.bad) {
        int sym = huff_decode(&s, &code);
        if (sym < 0)
          break;
        if (sym < 16) {
          lengths[i++] = (uint8_t)sym;
        } else {
          int rep, val = 0;
          if (sym == 16) {
            if (i == 0) {
              s.bad = 1;
              break;
            }
            val = lengths[i - 1], rep = 3 + (int)take_bits(&s, 2);
          } else if (sym == 17) {
            rep = 3 + (int)take_bits(&s, 3);
          } else {
            rep = 11 + (int)take_bits(&s, 7);
          }
          if (i + rep > n_lit + n_dist) {
            s.bad = 1;
            break;
          }
          while (rep--)
            lengths[i++] = (uint8_t)val;
        }
      }
      if (!s.bad) {
        Huff lit, dist;
        huff_build(&lit, lengths, n_lit);
        huff_build(&dist, lengths + n_lit, n_dist);
        inflate_codes(&s, &lit, &dist);
      }
    } else {
      s.bad = 1;
    }
  } while (!last && !s.bad);
  if (s.bad) {
    free(s.out);
    return fail_null("PNG: corrupt deflate stream");
  }
  *n_out = s.n_out;
  return s.out;
}

/* ================================================================== PNG ========================================== */

static int paeth(int a, int b, int c) {
  int p = a + b - c, pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
  return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}
/* reverses the per-scanline filters of one (sub)image in place; `raw` holds h rows of 1 + stride bytes */
static int png_unfilter(uint8_t *raw, size_t stride, int h, int bpp) {
  for (int y = 0; y < h; y++) {
    uint8_t *row = raw + (size_t)y * (stride + 1);
    const uint8_t *up = y ? row - stride : NULL; /* previous row's DATA (its filter byte sits before it) */
    int filter = row[0];
    row++;
    if (filter > 4)
      return -1;
    for (size_t x = 0; x < stride; x++) {
      int a = x >= (size_t)bpp ? row[x - bpp] : 0, b = up ? up[x] : 0, c = (up && x >= (size_t)bpp) ? up[x - bpp] : 0;
      int add = filter == 1 ? a : filter == 2 ? b : filter == 3 ? (a + b) >> 1 : filter == 4 ? paeth(a, b, c) : 0;
      row[x] = (uint8_t)(row[x] + add);
    }
  }
  return 0;
}

static uint8_t *png_decode(const uint8_t *data, size_t n, int *width, int *height) {
  if (n < 8 + 25)
    return fail_null("PNG: truncated");
  size_t at = 8;
  uint32_t w = 0, h = 0;
  int depth = 0, color = 0, interlace = 0, have_header = 0, n_pal = 0;
  uint8_t palette[256 * 3];
  uint8_t *idat = NULL;
  size_t n_idat = 0;
  for (;;) {
    if (at + 12 > n) {
      free(idat);
      return fail_null("PNG: truncated chunk");
    }
    uint32_t len = be32(data + at);
    const uint8_t *type = data + at + 4, *body = data + at + 8;
    if (len > n - at - 12) {
      free(idat);
      return fail_null("PNG: chunk runs past the end of the file");
    }
    if (memcmp(type, "IHDR", 4) == 0 && len == 13) {
      w = be32(body), h = be32(body + 4);
      depth = body[8], color = body[9], interlace = body[12];
      have_header = 1;
      if (w == 0 || h == 0 || w > (1u << 24) || h > (1u << 24) || body[10] != 0 || body[11] != 0 || interlace > 1)
        return fail_null("PNG: unsupported header");
      if (!(depth == 1 || depth == 2 || depth == 4 || depth == 8 || depth == 16) ||
          !(color == 0 || color == 2 || color == 3 || color == 4 || color == 6) || (color == 3 && depth == 16) ||
          ((color == 2 || color == 4 || color == 6) && depth < 8))
        return fail_null("PNG: unsupported colour type / bit depth");
    } else if (memcmp(type, "PLTE", 4) == 0) {
      if (len > 768 || len % 3) {
        free(idat);
        return fail_null("PNG: bad palette");
      }
      n_pal = (int)(len / 3);
      memcpy(palette, body, len);
    } else if (memcmp(type, "IDAT", 4) == 0) {
      uint8_t *grown = realloc(idat, n_idat + len + 1);
      if (grown == NULL) {
        free(idat);
        return fail_null("PNG: out of memory");
      }
      idat = grown;
      memcpy(idat + n_idat, body, len);
      n_idat += len;
    } else if (memcmp(type, "IEND", 4) == 0) {
      break;
    }
    at += 12 + (size_t)len;
  }
  if (!have_header || idat == NULL || (color == 3 && n_pal == 0)) {
    free(idat);
    return fail_null("PNG: missing IHDR, IDAT or PLTE");
  }
  size_t n_raw = 0;
  uint8_t *raw = zlib_inflate(idat, n_idat, &n_raw);
  free(idat);
  if (raw == NULL)
    return NULL;

  const int channels = color == 0 ? 1 : color == 2 ? 3 : color == 3 ? 1 : color == 4 ? 2 : 4;
  const int bits_pp = channels * depth, bpp = bits_pp >= 8 ? bits_pp / 8 : 1;
  uint8_t *rgb = my_malloc((size_t)w * h * 3);
  /* Adam7 passes (one pass covering everything when not interlaced) */
  static const int x0[7] = {0, 4, 0, 2, 0, 1, 0}, y0[7] = {0, 0, 4, 0, 2, 0, 1}, dx[7] = {8, 8, 4, 4, 2, 2, 1},
                   dy[7] = {8, 8, 8, 4, 4, 2, 2};
  size_t used = 0;
  const int n_pass = interlace ? 7 : 1;
  for (int p = 0; p < n_pass; p++) {
    const int px0 = interlace ? x0[p] : 0, py0 = interlace ? y0[p] : 0, pdx = interlace ? dx[p] : 1,
              pdy = interlace ? dy[p] : 1;
    const int pw = ((int)w - px0 + pdx - 1) / pdx, ph = ((int)h - py0 + pdy - 1) / pdy;
    if (pw <= 0 || ph <= 0)
      continue;
    const size_t stride = ((size_t)pw * bits_pp + 7) / 8;
    if (used + (stride + 1) * (size_t)ph > n_raw || png_unfilter(raw + used, stride, ph, bpp) != 0) {
      free(raw), free(rgb);
      return fail_null("PNG: image data too short or bad filter");
    }
    for (int y = 0; y < ph; y++) {
      const uint8_t *row = raw + used + (size_t)y * (stride + 1) + 1;
      for (int x = 0; x < pw; x++) {
        uint8_t s[4] = {0, 0, 0, 0}; /* the pixel's samples as 8-bit values */
        if (depth == 8) {
          for (int c = 0; c < channels; c++)
            s[c] = row[(size_t)x * channels + c];
        } else if (depth == 16) { /* stbi_load (8-bit API) keeps the high byte of each 16-bit sample */
          for (int c = 0; c < channels; c++)
            s[c] = row[((size_t)x * channels + c) * 2];
        } else { /* 1, 2, 4 bits: gray (scaled to 0..255) or a palette index (not scaled) */
          const int per_byte = 8 / depth, shift = (per_byte - 1 - x % per_byte) * depth;
          const int v = (row[x / per_byte] >> shift) & ((1 << depth) - 1);
          s[0] = (uint8_t)(color == 3 ? v : v * (depth == 1 ? 255 : depth == 2 ? 85 : 17));
        }
        uint8_t *out = rgb + ((size_t)(py0 + y * pdy) * w + (size_t)(px0 + x * pdx)) * 3;
        if (color == 3) {
          if (s[0] >= n_pal) { /* stb reads past the entries it was given (zeroed table): black */
            out[0] = out[1] = out[2] = 0;
          } else {
            memcpy(out, palette + 3 * s[0], 3);
          }
        } else if (channels <= 2) {
          out[0] = out[1] = out[2] = s[0]; /* gray (alpha dropped) */
        } else {
          out[0] = s[0], out[1] = s[1], out[2] = s[2]; /* RGB (alpha dropped, not composited) */
        }
      }
    }
    used += (stride + 1) * (size_t)ph;
  }
  free(raw);
  *width = (int)w, *height = (int)h;
  return rgb;
}

/* ================================================================== baseline JPEG ================================ */

typedef struct {
  uint8_t size[257]; /* code length of the k-th code, in code order; 0-terminated */
  uint16_t code[256];
  uint8_t value[256];
  int n;
  /* decoding by length: the largest code of each length (left-aligned to 16 bits, exclusive) and where its codes start */
  uint32_t max_code[18];
  int delta[17];
} JHuff;

typedef struct {
  int id, h, v, tq, td, ta; /* sampling factors, quantisation table, DC / AC Huffman tables */
  int w_blocks, h_blocks;   /* blocks per row / column of the padded plane */
  int x, y;                 /* size in samples: ceil(image * h / h_max) */
  int stride;               /* bytes per row of the padded plane */
  uint8_t *plane;
  int dc_pred;
} JComp;

typedef struct {
  const uint8_t *data;
  size_t n, at;
  uint32_t bits; /* MSB-first bit buffer */
  int n_bits;
  int marker; /* a marker met inside the entropy-coded data (0xFF..), 0 when none */
  uint16_t quant[4][64];
  JHuff dc[4], ac[4];
  int have_dc[4], have_ac[4];
  JComp comp[4];
  int n_comp, width, height, h_max, v_max, restart_interval, adobe_transform, jfif;
  int bad;
} Jpeg;

static const uint8_t kZigzag[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                                    41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                                    30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

static int jhuff_build(JHuff *h, const uint8_t counts[16], const uint8_t *values, int n_values) {
  int k = 0;
  for (int len = 1; len <= 16; len++)
    for (int i = 0; i < counts[len - 1]; i++) {
      if (k >= 256)
        return -1;
      h->size[k++] = (uint8_t)len;
    }
  h->size[k] = 0;
  h->n = k;
  if (k != n_values)
    return -1;
  memcpy(h->value, values, (size_t)k);
  uint32_t code = 0;
  k = 0;
  for (int len = 1; len <= 16; len++) {
    h->delta[len] = k - (int)code;
    while (h->size[k] == len)
      h->code[k++] = (uint16_t)code++;
    if (code > (1u << len))
      return -1;
    h->max_code[len] = code << (16 - len);
    code <<= 1;
  }
  h->max_code[17] = 0xffffffffu;
  return 0;
}

static void jpeg_fill(Jpeg *j) {
  while (j->n_bits <= 24) {
    uint32_t b = 0;
    if (!j->marker && j->at < j->n) {
      b = j->data[j->at++];
      if (b == 0xff) {
        uint32_t c = j->at < j->n ? j->data[j->at++] : 0xd9;
        while (c == 0xff)
          c = j->at < j->n ? j->data[j->at++] : 0xd9; /* fill bytes */
        if (c != 0) { /* a marker ends the entropy-coded segment: zeros from here on */
          j->marker = (int)c;
          b = 0;
        }
      }
    }
    j->bits |= b << (24 - j->n_bits);
    j->n_bits += 8;
  }
}
static int jpeg_bits(Jpeg *j, int n) {
  if (n == 0)
    return 0;
  if (j->n_bits < n)
    jpeg_fill(j);
  uint32_t v = j->bits >> (32 - n);
  j->bits <<= n;
  j->n_bits -= n;
  return (int)v;
}
static int jpeg_symbol(Jpeg *j, const JHuff *h) {
  if (j->n_bits < 16)
    jpeg_fill(j);
  const uint32_t top = j->bits >> 16;
  int len = 1;
  while (len <= 16 && top >= h->max_code[len])
    len++;
  if (len > 16 || j->n_bits < len) {
    j->bad = 1;
    return 0;
  }
  const int k = (int)(j->bits >> (32 - len)) + h->delta[len];
  j->bits <<= len;
  j->n_bits -= len;
  if (k < 0 || k >= h->n) {
    j->bad = 1;
    return 0;
  }
  return h->value[k];
}
/* the `n`-bit two's-complement-like extension of JPEG (F.2.2.1) */
static int jpeg_extend(Jpeg *j, int n) {
  if (n == 0)
    return 0;
  const int v = jpeg_bits(j, n);
  return v < (1 << (n - 1)) ? v - (1 << n) + 1 : v;
}

/* One pass of the inverse DCT over 8 values (IJG jidctint "islow", constants scaled by 4096): the even part in e[0..3],
 * the odd part in o[0..3]; outputs are e[k] + o[3-k] (first four) and e[k] - o[3-k] (mirrored last four). */
#define FIX(x) ((int)((x) * 4096 + 0.5))
static void idct_pass(int s0, int s1, int s2, int s3, int s4, int s5, int s6, int s7, int e[4], int o[4]) {
  int p1 = (s2 + s6) * FIX(0.5411961f);
  int t2 = p1 + s6 * FIX(-1.847759065f);
  int t3 = p1 + s2 * FIX(0.765366865f);
  int t0 = (s0 + s4) * 4096, t1 = (s0 - s4) * 4096;
  e[0] = t0 + t3, e[3] = t0 - t3, e[1] = t1 + t2, e[2] = t1 - t2;
  int q0 = s7, q1 = s5, q2 = s3, q3 = s1;
  int p3 = q0 + q2, p4 = q1 + q3, pa = q0 + q3, pb = q1 + q2;
  int p5 = (p3 + p4) * FIX(1.175875602f);
  q0 = q0 * FIX(0.298631336f), q1 = q1 * FIX(2.053119869f), q2 = q2 * FIX(3.072711026f), q3 = q3 * FIX(1.501321110f);
  pa = p5 + pa * FIX(-0.899976223f), pb = p5 + pb * FIX(-2.562915447f);
  p3 = p3 * FIX(-1.961570560f), p4 = p4 * FIX(-0.390180644f);
  o[3] = q3 + pa + p4, o[2] = q2 + pb + p3, o[1] = q1 + pb + p4, o[0] = q0 + pa + p3;
}
static uint8_t clamp_u8(int v) { return (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v); }
static void idct_block(uint8_t *out, int stride, const short d[64]) {
  int mid[64], e[4], o[4];
  for (int c = 0; c < 8; c++) { /* columns: keep two extra bits */
    idct_pass(d[c], d[8 + c], d[16 + c], d[24 + c], d[32 + c], d[40 + c], d[48 + c], d[56 + c], e, o);
    for (int k = 0; k < 4; k++) {
      mid[8 * k + c] = (e[k] + 512 + o[3 - k]) >> 10;
      mid[8 * (7 - k) + c] = (e[k] + 512 - o[3 - k]) >> 10;
    }
  }
  for (int r = 0; r < 8; r++) { /* rows: drop 17 bits, rounded, with the +128 level shift folded in */
    const int *m = mid + 8 * r;
    idct_pass(m[0], m[1], m[2], m[3], m[4], m[5], m[6], m[7], e, o);
    uint8_t *row = out + (size_t)r * stride;
    for (int k = 0; k < 4; k++) {
      const int base = e[k] + 65536 + (128 << 17);
      row[k] = clamp_u8((base + o[3 - k]) >> 17);
      row[7 - k] = clamp_u8((base - o[3 - k]) >> 17);
    }
  }
}

static void jpeg_block(Jpeg *j, JComp *c, int bx, int by) {
  short coef[64];
  memset(coef, 0, sizeof(coef));
  const uint16_t *q = j->quant[c->tq];
  int t = jpeg_symbol(j, &j->dc[c->td]);
  if (t > 15) {
    j->bad = 1;
    return;
  }
  c->dc_pred += jpeg_extend(j, t);
  coef[0] = (short)(c->dc_pred * q[0]);
  for (int k = 1; k < 64 && !j->bad;) {
    const int rs = jpeg_symbol(j, &j->ac[c->ta]), r = rs >> 4, s = rs & 15;
    if (s == 0) {
      if (rs != 0xf0)
        break; /* end of block */
      k += 16;
    } else {
      k += r;
      if (k > 63) {
        j->bad = 1;
        return;
      }
      const int z = kZigzag[k++];
      coef[z] = (short)(jpeg_extend(j, s) * q[z]);
    }
  }
  idct_block(c->plane + ((size_t)by * 8) * c->stride + (size_t)bx * 8, c->stride, coef);
}

static void jpeg_restart(Jpeg *j) {
  j->bits = 0, j->n_bits = 0;
  if (!j->marker) { /* the RSTn marker has not been met by the bit reader yet: find it */
    while (j->at + 1 < j->n && !(j->data[j->at] == 0xff && j->data[j->at + 1] >= 0xd0 && j->data[j->at + 1] <= 0xd7))
      j->at++;
    if (j->at + 1 < j->n)
      j->at += 2;
  }
  j->marker = 0;
  for (int i = 0; i < j->n_comp; i++)
    j->comp[i].dc_pred = 0;
}

static int jpeg_scan(Jpeg *j, const int *order, int n_scan) {
  j->bits = 0, j->n_bits = 0, j->marker = 0;
  for (int i = 0; i < j->n_comp; i++)
    j->comp[i].dc_pred = 0;
  int todo = j->restart_interval ? j->restart_interval : 0x7fffffff;
  if (n_scan == 1) { /* one component: its own blocks in raster order, only those that hold image samples */
    JComp *c = &j->comp[order[0]];
    const int w = (c->x + 7) >> 3, h = (c->y + 7) >> 3;
    for (int by = 0; by < h; by++)
      for (int bx = 0; bx < w; bx++) {
        jpeg_block(j, c, bx, by);
        if (j->bad)
          return -1;
        if (--todo <= 0) {
          if (!(by == h - 1 && bx == w - 1))
            jpeg_restart(j);
          todo = j->restart_interval;
        }
      }
  } else { /* interleaved: MCU by MCU, h x v blocks of every component */
    const int mcu_x = (j->width + 8 * j->h_max - 1) / (8 * j->h_max), mcu_y = (j->height + 8 * j->v_max - 1) / (8 * j->v_max);
    for (int my = 0; my < mcu_y; my++)
      for (int mx = 0; mx < mcu_x; mx++) {
        for (int k = 0; k < n_scan; k++) {
          JComp *c = &j->comp[order[k]];
          for (int v = 0; v < c->v; v++)
            for (int h = 0; h < c->h; h++) {
              jpeg_block(j, c, mx * c->h + h, my * c->v + v);
              if (j->bad)
                return -1;
            }
        }
        if (--todo <= 0) {
          if (!(my == mcu_y - 1 && mx == mcu_x - 1))
            jpeg_restart(j);
          todo = j->restart_interval;
        }
      }
  }
  return 0;
}

/* chroma plane -> full resolution row `row` (0-based, full-resolution) into `out` (width samples).  JFIF-centred
 * "fancy" upsampling for factors of 2 (3:1 taps), nearest neighbour for anything else. */
static void upsample_row(const JComp *c, int hs, int vs, int row, int width, uint8_t *out, uint8_t *tmp) {
  const int wl = (width + hs - 1) / hs; /* low-resolution samples per row that matter */
  /* which two low-resolution rows, and which is the nearer one */
  const uint8_t *near_row, *far_row;
  if (vs == 1) {
    near_row = far_row = c->plane + (size_t)row * c->stride;
  } else if (vs == 2) {
    const int r = row >> 1, other = (row & 1) ? r + 1 : r - 1;
    const int last = c->y - 1;
    near_row = c->plane + (size_t)(r > last ? last : r) * c->stride;
    far_row = c->plane + (size_t)(other < 0 ? 0 : other > last ? last : other) * c->stride;
  } else {
    const int r = row / vs;
    near_row = far_row = c->plane + (size_t)(r > c->y - 1 ? c->y - 1 : r) * c->stride;
  }
  if (hs == 1 && vs == 1) {
    memcpy(out, near_row, (size_t)width);
  } else if (hs == 1 && vs == 2) {
    for (int i = 0; i < wl; i++)
      out[i] = (uint8_t)((3 * near_row[i] + far_row[i] + 2) >> 2);
  } else if (hs == 2 && vs == 1) {
    const uint8_t *in = near_row;
    if (wl == 1) {
      tmp[0] = tmp[1] = in[0];
    } else {
      tmp[0] = in[0];
      tmp[1] = (uint8_t)((in[0] * 3 + in[1] + 2) >> 2);
      int i;
      for (i = 1; i < wl - 1; i++) {
        const int n3 = 3 * in[i] + 2;
        tmp[2 * i] = (uint8_t)((n3 + in[i - 1]) >> 2);
        tmp[2 * i + 1] = (uint8_t)((n3 + in[i + 1]) >> 2);
      }
      tmp[2 * i] = (uint8_t)((in[wl - 2] * 3 + in[wl - 1] + 2) >> 2);
      tmp[2 * i + 1] = in[wl - 1];
    }
    memcpy(out, tmp, (size_t)width);
  } else if (hs == 2 && vs == 2) {
    if (wl == 1) {
      tmp[0] = tmp[1] = (uint8_t)((3 * near_row[0] + far_row[0] + 2) >> 2);
    } else {
      int t1 = 3 * near_row[0] + far_row[0];
      tmp[0] = (uint8_t)((t1 + 2) >> 2);
      for (int i = 1; i < wl; i++) {
        const int t0 = t1;
        t1 = 3 * near_row[i] + far_row[i];
        tmp[2 * i - 1] = (uint8_t)((3 * t0 + t1 + 8) >> 4);
        tmp[2 * i] = (uint8_t)((3 * t1 + t0 + 8) >> 4);
      }
      tmp[2 * wl - 1] = (uint8_t)((t1 + 2) >> 2);
    }
    memcpy(out, tmp, (size_t)width);
  } else {
    for (int i = 0; i < width; i++)
      out[i] = near_row[i / hs];
  }
}

#define YCC_FIX(x) (((int)((x) * 4096.0f + 0.5f)) << 8)
static uint8_t *jpeg_decode(const uint8_t *data, size_t n, int *width, int *height) {
  Jpeg *j = calloc(1, sizeof(Jpeg));
  if (j == NULL)
    return fail_null("JPEG: out of memory");
  j->data = data, j->n = n, j->at = 2;
  j->adobe_transform = -1;
  uint8_t *rgb = NULL;
  const char *why = NULL;
  int have_frame = 0, done = 0;
  while (!done && why == NULL) {
    /* next marker */
    int m = j->marker;
    j->marker = 0;
    if (!m) {
      while (j->at < j->n && j->data[j->at] != 0xff)
        j->at++;
      while (j->at < j->n && j->data[j->at] == 0xff)
        j->at++;
      if (j->at >= j->n) {
        why = "JPEG: no end-of-image marker";
        break;
      }
      m = j->data[j->at++];
    }
    if (m == 0xd9) {
      done = 1;
      break;
    }
    if (m == 0x01 || (m >= 0xd0 && m <= 0xd7))
      continue;
    if (j->at + 2 > j->n) {
      why = "JPEG: truncated";
      break;
    }
    const size_t len = be16(j->data + j->at);
    if (len < 2 || j->at + len > j->n) {
      why = "JPEG: bad segment length";
      break;
    }
    const uint8_t *seg = j->data + j->at + 2, *seg_end = j->data + j->at + len;
    j->at += len;
    if (m == 0xdb) { /* quantisation tables */
      while (seg < seg_end) {
        const int pq = seg[0] >> 4, tq = seg[0] & 15;
        seg++;
        if (tq > 3 || pq > 1 || seg + (pq ? 128 : 64) > seg_end) {
          why = "JPEG: bad quantisation table";
          break;
        }
        for (int i = 0; i < 64; i++) {
          j->quant[tq][kZigzag[i]] = (uint16_t)(pq ? be16(seg + 2 * i) : seg[i]);
        }
        seg += pq ? 128 : 64;
      }
    } else if (m == 0xc4) { /* Huffman tables */
      while (seg < seg_end && why == NULL) {
        const int tc = seg[0] >> 4, th = seg[0] & 15;
        if (tc > 1 || th > 3 || seg + 17 > seg_end) {
          why = "JPEG: bad Huffman table header";
          break;
        }
        int total = 0;
        for (int i = 0; i < 16; i++)
          total += seg[1 + i];
        if (total > 256 || seg + 17 + total > seg_end || jhuff_build(tc ? &j->ac[th] : &j->dc[th], seg + 1, seg + 17, total)) {
          why = "JPEG: bad Huffman table";
          break;
        }
        (tc ? j->have_ac : j->have_dc)[th] = 1;
        seg += 17 + total;
      }
    } else if (m == 0xc0 || m == 0xc1) { /* baseline / extended sequential, Huffman */
      if (have_frame || len < 8 || seg[0] != 8) {
        why = seg[0] != 8 ? "JPEG: only 8-bit samples are decoded" : "JPEG: bad frame header";
        break;
      }
      j->height = (int)be16(seg + 1), j->width = (int)be16(seg + 3), j->n_comp = seg[5];
      if (j->height == 0 || j->width == 0 || !(j->n_comp == 1 || j->n_comp == 3) || len < 8 + 3 * (size_t)j->n_comp) {
        why = j->n_comp == 4 ? "JPEG: CMYK / YCCK images are not decoded" : "JPEG: bad frame header";
        break;
      }
      for (int i = 0; i < j->n_comp; i++) {
        JComp *c = &j->comp[i];
        c->id = seg[6 + 3 * i], c->h = seg[7 + 3 * i] >> 4, c->v = seg[7 + 3 * i] & 15, c->tq = seg[8 + 3 * i];
        if (c->h < 1 || c->h > 4 || c->v < 1 || c->v > 4 || c->tq > 3) {
          why = "JPEG: bad sampling factors";
          break;
        }
        j->h_max = c->h > j->h_max ? c->h : j->h_max, j->v_max = c->v > j->v_max ? c->v : j->v_max;
      }
      if (why)
        break;
      const int mcu_x = (j->width + 8 * j->h_max - 1) / (8 * j->h_max), mcu_y = (j->height + 8 * j->v_max - 1) / (8 * j->v_max);
      for (int i = 0; i < j->n_comp; i++) {
        JComp *c = &j->comp[i];
        if (j->h_max % c->h || j->v_max % c->v) {
          why = "JPEG: fractional sampling ratios are not decoded";
          break;
        }
        c->x = (j->width * c->h + j->h_max - 1) / j->h_max, c->y = (j->height * c->v + j->v_max - 1) / j->v_max;
        c->w_blocks = mcu_x * c->h, c->h_blocks = mcu_y * c->v;
        c->stride = c->w_blocks * 8;
        c->plane = calloc((size_t)c->stride * c->h_blocks * 8 + 16, 1);
        if (c->plane == NULL) {
          why = "JPEG: out of memory";
          break;
        }
      }
      have_frame = 1;
    } else if (m == 0xc2 || (m >= 0xc5 && m <= 0xcf && m != 0xc8 && m != 0xcc)) {
      why = m == 0xc2 ? "JPEG: progressive images are not decoded (re-save as baseline)"
                      : "JPEG: lossless / arithmetic-coded images are not decoded";
    } else if (m == 0xdd) {
      if (len != 4) {
        why = "JPEG: bad restart interval";
        break;
      }
      j->restart_interval = (int)be16(seg);
    } else if (m == 0xe0) { /* JFIF: the components are Y, Cb, Cr by definition */
      if (len >= 7 && memcmp(seg, "JFIF", 5) == 0)
        j->jfif = 1;
    } else if (m == 0xee) { /* Adobe: colour transform flag */
      if (len >= 14 && memcmp(seg, "Adobe", 5) == 0)
        j->adobe_transform = seg[11];
    } else if (m == 0xda) { /* scan */
      if (!have_frame) {
        why = "JPEG: scan before frame header";
        break;
      }
      const int n_scan = seg[0];
      int order[4];
      if (n_scan < 1 || n_scan > j->n_comp || len != 6 + 2 * (size_t)n_scan) {
        why = "JPEG: bad scan header";
        break;
      }
      for (int k = 0; k < n_scan && why == NULL; k++) {
        int which = -1;
        for (int i = 0; i < j->n_comp; i++)
          if (j->comp[i].id == seg[1 + 2 * k])
            which = i;
        if (which < 0) {
          why = "JPEG: scan names an unknown component";
          break;
        }
        order[k] = which;
        j->comp[which].td = seg[2 + 2 * k] >> 4, j->comp[which].ta = seg[2 + 2 * k] & 15;
        if (j->comp[which].td > 3 || j->comp[which].ta > 3 || !j->have_dc[j->comp[which].td] ||
            !j->have_ac[j->comp[which].ta])
          why = "JPEG: scan uses an undefined Huffman table";
      }
      if (why)
        break;
      if (seg[1 + 2 * n_scan] != 0 || seg[2 + 2 * n_scan] != 63 || seg[3 + 2 * n_scan] != 0) {
        why = "JPEG: bad spectral selection for a sequential scan";
        break;
      }
      if (jpeg_scan(j, order, n_scan) != 0)
        why = "JPEG: corrupt entropy-coded data";
    }
    /* every other segment (APPn, COM, ...) is skipped */
  }
  if (why == NULL && !have_frame)
    why = "JPEG: no frame";
  if (why == NULL) {
    const int w = j->width, h = j->height;
    rgb = my_malloc((size_t)w * h * 3);
    uint8_t *line[3], *tmp = my_malloc((size_t)w + 16 + 8 * (size_t)j->h_max);
    for (int i = 0; i < 3; i++)
      line[i] = my_malloc((size_t)w + 16);
    /* three components are Y, Cb, Cr unless an Adobe segment says "no transform" (or names them R, G, B) */
    const int is_rgb = j->n_comp == 3 && ((j->adobe_transform == 0 && !j->jfif) ||
                                          (j->comp[0].id == 'R' && j->comp[1].id == 'G' && j->comp[2].id == 'B'));
    for (int y = 0; y < h; y++) {
      for (int i = 0; i < j->n_comp; i++)
        upsample_row(&j->comp[i], j->h_max / j->comp[i].h, j->v_max / j->comp[i].v, y, w, line[i], tmp);
      uint8_t *out = rgb + (size_t)y * w * 3;
      if (j->n_comp == 1) {
        for (int x = 0; x < w; x++)
          out[3 * x] = out[3 * x + 1] = out[3 * x + 2] = line[0][x];
      } else if (is_rgb) {
        for (int x = 0; x < w; x++)
          out[3 * x] = line[0][x], out[3 * x + 1] = line[1][x], out[3 * x + 2] = line[2][x];
      } else {
        for (int x = 0; x < w; x++) { /* 20-bit fixed point, the chroma-to-green term of Cb truncated to 16 bits */
          const int yf = (line[0][x] << 20) + (1 << 19), cb = line[1][x] - 128, cr = line[2][x] - 128;
          int r = yf + cr * YCC_FIX(1.40200f);
          int g = yf + (cr * -YCC_FIX(0.71414f)) + (int)((unsigned)(cb * -YCC_FIX(0.34414f)) & 0xffff0000u);
          int b = yf + cb * YCC_FIX(1.77200f);
          out[3 * x] = clamp_u8(r >> 20), out[3 * x + 1] = clamp_u8(g >> 20), out[3 * x + 2] = clamp_u8(b >> 20);
        }
      }
    }
    for (int i = 0; i < 3; i++)
      free(line[i]);
    free(tmp);
    *width = w, *height = h;
  }
  for (int i = 0; i < 4; i++)
    free(j->comp[i].plane);
  free(j);
  if (why != NULL)
    return fail_null(why);
  return rgb;
}

/* ================================================================== entry point ================================== */

/* Decodes `filename` (binary PPM, PNG or baseline JPEG, recognised by content) to a malloc'ed RGB8 buffer.
 * Returns NULL with *exists = 0 when the file cannot be opened, NULL with *exists = 1 and a reason in
 * rtc_image_error() when it is there but cannot be decoded. */
uint8_t *rtc_load_image(const char *filename, int *width, int *height, int *exists) {
  g_why = "";
  *exists = 0;
  FILE *f = fopen(filename, "rb");
  if (f == NULL)
    return fail_null("cannot open the file");
  *exists = 1;
  fseek(f, 0, SEEK_END);
  long size = ftell(f);
  fseek(f, 0, SEEK_SET);
  uint8_t *data = size > 0 ? malloc((size_t)size) : NULL;
  if (data == NULL || fread(data, 1, (size_t)size, f) != (size_t)size) {
    fclose(f);
    free(data);
    return fail_null("cannot read the file");
  }
  fclose(f);
  uint8_t *rgb = NULL;
  static const uint8_t png_magic[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
  if (size >= 8 && memcmp(data, png_magic, 8) == 0) {
    rgb = png_decode(data, (size_t)size, width, height);
  } else if (size >= 4 && data[0] == 0xff && data[1] == 0xd8) {
    rgb = jpeg_decode(data, (size_t)size, width, height);
  } else if (size >= 2 && data[0] == 'P' && data[1] == '6') {
    rgb = rtc_load_ppm(filename, width, height);
    if (rgb == NULL)
      g_why = "PPM: only binary P6 with maxval 255 is read";
  } else {
    g_why = "not a PNG, JPEG or binary PPM file";
  }
  free(data);
  return rgb;
}

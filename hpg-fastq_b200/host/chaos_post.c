/*
 * chaos_post.c -- what the reference does with the chaos-game tables after the reads are counted
 * (SURVEY.md 8f rank 3): genomic-signature loader, normalised difference matrix with its range / mean / standard
 * deviation, and the three binary PGM (P5) images.  Mirrors, on flat row-major arrays taken from the counter block
 * (HPGQ_SEC_CG_SEQ / HPGQ_SEC_CG_Q, filled on the GPU), the reference's old/chaos_game.c:
 *   chaos_game_load_table_gs_direct   :298-320     hq_cg_load_signature
 *   chaos_game_calculate_table_dif    :322-377     hq_cg_diff
 *   chaos_game_validate_table_dif     :379-405     hq_cg_diff (mean, sd; the reference's accumulators start from
 *                                                  indeterminate values, here from 0)
 *   chaos_game_write_table_images     :407-465     hq_cg_write_images
 *   chaos_game_normalize_quality_table_ :479-494, chaos_game_absolute_diff_table_ :496-512,
 *   chaos_game_generate_pgm_file_     :514-593     hq_cg_pixels_*, hq_cg_write_pgm
 * The reference keeps 32-bit cells; the u64 counters of this implementation are narrowed to 32 bits first, so the
 * images agree with the reference wherever the reference itself does not overflow.  Pixel arithmetic follows the
 * reference's types step by step (u32 -> float, float * double rounded to float, truncation to int, wrap to uchar),
 * which is what makes the files byte-identical.  tests/test_chaos_post.py checks every function against the
 * reference's own code (oracle/_ref).
 */
#include "hpgq_host.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#define CG_MIN_K_IMAGE 7       /* MIN_K_IMAGE_VALUE, old/chaos_game.h:45 */
#define CG_MIN_IMAGE_PIXELS 128 /* MIN_IMAGE_PIXEL_SIZE, :46 */
#define CG_MAX_QUALITY 62      /* MAX_QUALITY_IN_TABLE, :40 */
#define CG_GS_HEADER_BYTES 196 /* sizeof(header_gs_t): char[180] + 4 x unsigned, old/chaos_game.h:65-70 */

/* one cell -> one grey value, old/chaos_game.c:537-542 */
static inline uint8_t cg_pixel(uint32_t cell, double norm) {
  float v = (float)cell;
  v = (float)((double)v * norm);
  const int iv = (int)v;
  return (uint8_t)iv;
}

void hq_cg_pixels_seq(const uint64_t *table_seq, uint32_t dim, uint64_t fq_word_count, uint8_t *pix) {
  /* fq_norm = 128 / (words / 4^k), old/chaos_game.c:432-439; the caller has checked fq_word_count > 0 */
  double norm = (double)(uint32_t)fq_word_count;
  norm = norm / (double)((uint64_t)dim * dim);
  norm = 128.0 / norm;
  for (uint64_t i = 0; i < (uint64_t)dim * dim; i++) pix[i] = cg_pixel((uint32_t)table_seq[i], norm);
}

void hq_cg_pixels_quality(const uint64_t *table_seq, const uint64_t *table_q, uint32_t dim, int k, uint8_t *pix) {
  /* quality sum / word size / word count in unsigned integer arithmetic (:479-494), then 256 / 62 (:455) */
  const double norm = 256.0 / CG_MAX_QUALITY;
  for (uint64_t i = 0; i < (uint64_t)dim * dim; i++) {
    uint32_t q = 0;
    const uint32_t n = (uint32_t)table_seq[i];
    if (n > 0) q = ((uint32_t)table_q[i] / (uint32_t)k) / n;
    pix[i] = cg_pixel(q, norm);
  }
}

int hq_cg_load_signature(const char *path, uint32_t dim, uint32_t *table_gs, uint32_t *ref_word_count) {
  FILE *f = fopen(path, "rb");
  if (!f) return -1;
  uint8_t hdr[CG_GS_HEADER_BYTES];
  int rc = -1;
  if (fread(hdr, 1, sizeof hdr, f) == sizeof hdr) {
    uint32_t words;
    memcpy(&words, hdr + 180 + 12, 4); /* word_size_k, dim_x, dim_y, ref_word_count */
    *ref_word_count = words;
    /* the reference reads dim rows of dim ints and ignores the header's own dimensions (:314-316) */
    if (fread(table_gs, sizeof(uint32_t), (size_t)dim * dim, f) == (size_t)dim * dim) rc = 0;
  }
  fclose(f);
  return rc;
}

int hq_cg_diff(const uint64_t *table_seq, const uint32_t *table_gs, uint32_t dim, uint64_t fq_word_count,
               uint32_t ref_word_count, int32_t *dif, hq_cg_diff_stats_t *st) {
  const double cells = (double)((uint64_t)dim * dim);
  double fq_norm = (double)(1.0 * (uint32_t)fq_word_count / cells), gs_norm = (double)(1.0 * ref_word_count / cells);
  if (!(fq_norm > 0.0) || !(gs_norm > 0.0)) return -1; /* LOG_FATAL in the reference, :334-337,346-349 */
  fq_norm = 128.0 / fq_norm;
  gs_norm = 128.0 / gs_norm;
  int highest = -32768, lowest = 32768;
  for (uint64_t i = 0; i < (uint64_t)dim * dim; i++) {
    const int v = (int)(((uint32_t)table_seq[i] * fq_norm) - (table_gs[i] * gs_norm));
    dif[i] = v;
    if (highest < v) highest = v;
    if (lowest > v) lowest = v;
  }
  double mean = 0.0, sd = 0.0;
  for (uint64_t i = 0; i < (uint64_t)dim * dim; i++) mean += dif[i];
  mean /= (double)(1.0 * dim * dim);
  for (uint64_t i = 0; i < (uint64_t)dim * dim; i++) sd += pow(dif[i] - mean, 2);
  sd /= (double)(1.0 * dim * dim);
  sd = sqrt(sd);
  st->highest = highest;
  st->lowest = lowest;
  st->mean = mean;
  st->sd = sd;
  st->fq_norm = fq_norm;
  st->gs_norm = gs_norm;
  return 0;
}

void hq_cg_pixels_diff(const int32_t *dif, uint32_t dim, uint8_t *pix) {
  /* |d| clamped to 255 (:496-512), then through the common pixel path with norm 1 (:464) */
  for (uint64_t i = 0; i < (uint64_t)dim * dim; i++) {
    int v = abs(dif[i]);
    if (v > 255) v = 255;
    pix[i] = cg_pixel((uint32_t)(uint8_t)v, 1.0);
  }
}

int hq_cg_write_pgm(const char *path, const uint8_t *pix, uint32_t dim, int k) {
  FILE *f = fopen(path, "wb");
  if (!f) return -1;
  int rc = 0;
  if (k >= CG_MIN_K_IMAGE) {
    fprintf(f, "P5\n%d %d\n%d\n", (int)dim, (int)dim, 255);
    if (fwrite(pix, 1, (size_t)dim * dim, f) != (size_t)dim * dim) rc = -1;
  } else {
    /* small matrices are blown up to 128 x 128: every cell becomes a zoom x zoom square (:544-565) */
    const uint32_t zoom = 1u << (CG_MIN_K_IMAGE - k), out = CG_MIN_IMAGE_PIXELS;
    uint8_t *row = (uint8_t *)malloc(out);
    if (!row) {
      fclose(f);
      return -1;
    }
    fprintf(f, "P5\n%d %d\n%d\n", (int)out, (int)out, 255);
    for (uint32_t i = 0; i < out && !rc; i++) {
      for (uint32_t j = 0; j < out; j++) row[j] = pix[(uint64_t)(i / zoom) * dim + j / zoom];
      if (fwrite(row, 1, out, f) != out) rc = -1;
    }
    free(row);
  }
  if (fclose(f)) rc = -1;
  return rc;
}

int hq_cg_write_images(const hpgq_layout_t *lay, const uint64_t *block, const char *fq_path, const char *out_dir,
                       const char *signature_path, hq_cg_diff_stats_t *st_out) {
  const int k = (int)lay->cg_k;
  if (k <= 0) return 0;
  const uint32_t dim = 1u << k;
  const uint64_t cells = (uint64_t)dim * dim;
  const uint64_t *ts = block + lay->offset[HPGQ_SEC_CG_SEQ], *tq = block + lay->offset[HPGQ_SEC_CG_Q];
  const uint64_t words = block[HPGQ_S_CG_WORDS];
  if ((uint32_t)words == 0) return -2; /* "Error in fastq sequences normalization", old/chaos_game.c:438 */
  const char *slash = strrchr(fq_path, '/');
  const char *name = slash ? slash + 1 : fq_path;
  char base[4096], path[4200];
  snprintf(base, sizeof base, "%s/%s_k=%d", out_dir, name, k); /* K_VALUE_INFIX, old/chaos_game.h:49 */
  uint8_t *pix = (uint8_t *)malloc(cells);
  if (!pix) return -1;
  int rc = 0;
  hq_cg_pixels_seq(ts, dim, words, pix);
  snprintf(path, sizeof path, "%s_FG.pgm", base);
  if (hq_cg_write_pgm(path, pix, dim, k)) rc = -1;
  hq_cg_pixels_quality(ts, tq, dim, k, pix);
  snprintf(path, sizeof path, "%s_QQ.pgm", base);
  if (!rc && hq_cg_write_pgm(path, pix, dim, k)) rc = -1;
  if (!rc && signature_path) {
    uint32_t *gs = (uint32_t *)malloc(cells * sizeof(uint32_t));
    int32_t *dif = (int32_t *)malloc(cells * sizeof(int32_t));
    uint32_t ref_words = 0;
    hq_cg_diff_stats_t st;
    if (!gs || !dif || hq_cg_load_signature(signature_path, dim, gs, &ref_words) ||
        hq_cg_diff(ts, gs, dim, words, ref_words, dif, &st)) {
      rc = -3;
    } else {
      hq_cg_pixels_diff(dif, dim, pix);
      snprintf(path, sizeof path, "%s_FG_dif.pgm", base);
      if (hq_cg_write_pgm(path, pix, dim, k)) rc = -1;
      if (st_out) *st_out = st;
    }
    free(gs);
    free(dif);
  }
  free(pix);
  return rc;
}

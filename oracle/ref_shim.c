/*
 * ref_shim.c -- thin driver around the REFERENCE's own chaos_game.c (compiled from where it lies
 * under /root/reference/old by oracle/Makefile target `ref`; no reference source is copied here).
 * Test infrastructure only: lets tests/ call chaos_game_data_init + chaos_game_fill_tables
 * (old/chaos_game.c:97-130,165-267) on an SoA batch and read the tables back.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string.h>
#include "chaos_game.h" /* the reference header, found via -I/root/reference/old + oracle/ref_stubs */

int ref_chaos_run(int num_reads, char *seq, char *quality, int *data_indices, int *read_status, int mode, int k,
                  int base_quality, unsigned int *table_seq_out, unsigned int *table_q_out,
                  unsigned int *fq_word_count_out) {
  chaos_game_data_t *cg = (chaos_game_data_t *)calloc(1, sizeof(chaos_game_data_t));
  if (!cg) return -1;
  chaos_game_data_init(cg, NULL, k, 0, base_quality);
  fastq_batch_t fb;
  fb.num_reads = num_reads;
  fb.seq = seq;
  fb.quality = quality;
  fb.data_indices = data_indices;
  status_batch_t sb;
  sb.read_p = &fb;
  sb.read_status = read_status;
  chaos_game_fill_tables(cg, &sb, mode);
  const int dim = cg->dim_n;
  for (int x = 0; x < dim; ++x) {
    memcpy(table_seq_out + (size_t)x * dim, cg->table_seq[x], sizeof(unsigned int) * dim);
    memcpy(table_q_out + (size_t)x * dim, cg->table_q[x], sizeof(unsigned int) * dim);
  }
  *fq_word_count_out = cg->fq_word_count;
  cg->table_gs = NULL;
  chaos_game_data_free(cg);
  return 0;
}

/* file_utils helper of the missing common-libs, called by chaos_game_write_table_images
 * (old/chaos_game.c:416): copy the basename of `path` into `filename`. */
char *get_filename_from_path(char *path, char *filename) {
  const char *slash = strrchr(path, '/');
  strcpy(filename, slash ? slash + 1 : path);
  return filename;
}

/* Post-processing of the REFERENCE (old/chaos_game.c:298-465) on caller-provided tables: loads the signature file
 * with chaos_game_load_table_gs_direct, runs chaos_game_calculate_table_dif and chaos_game_write_table_images.
 * table_dif_out receives the difference matrix BEFORE write_table_images clamps it for the image. */
int ref_chaos_post(int k, const unsigned int *table_seq, const unsigned int *table_q, unsigned int fq_word_count,
                   char *gs_filename, char *fq_path, char *report_dir, int *table_dif_out, int *highest, int *lowest,
                   unsigned int *ref_word_count_out) {
  chaos_game_data_t *cg = (chaos_game_data_t *)calloc(1, sizeof(chaos_game_data_t));
  if (!cg) return -1;
  const int dim = 1 << k;
  unsigned int **gs = (unsigned int **)calloc(dim, sizeof(unsigned int *));
  for (int x = 0; x < dim; ++x) gs[x] = (unsigned int *)calloc(dim, sizeof(unsigned int));
  chaos_game_data_init(cg, gs, k, 0, 33);
  for (int x = 0; x < dim; ++x) {
    memcpy(cg->table_seq[x], table_seq + (size_t)x * dim, sizeof(unsigned int) * dim);
    memcpy(cg->table_q[x], table_q + (size_t)x * dim, sizeof(unsigned int) * dim);
  }
  cg->fq_word_count = fq_word_count;
  if (gs_filename) {
    cg->ref_word_count = chaos_game_load_table_gs_direct(gs, dim, gs_filename);
    *ref_word_count_out = cg->ref_word_count;
    chaos_game_calculate_table_dif(cg);
    for (int x = 0; x < dim; ++x) memcpy(table_dif_out + (size_t)x * dim, cg->table_dif[x], sizeof(int) * dim);
    *highest = cg->highest_table_dif_value;
    *lowest = cg->lowest_table_dif_value;
  }
  chaos_game_write_table_images(cg, fq_path, report_dir);
  fflush(NULL); /* chaos_game_generate_pgm_file_ never closes its FILE (old/chaos_game.c:514-593): the tail of each
                   image only reaches the file at exit */
  for (int x = 0; x < dim; ++x) free(gs[x]);
  free(gs);
  cg->table_gs = NULL;
  chaos_game_data_free(cg);
  return 0;
}

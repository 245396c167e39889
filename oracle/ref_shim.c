/*
 * ref_shim.c -- thin driver around the REFERENCE's own chaos_game.c (compiled from where it lies
 * under /root/reference/old by oracle/Makefile target `ref`; no reference source is copied here).
 * Test infrastructure only: lets tests/ call chaos_game_data_init + chaos_game_fill_tables
 * (old/chaos_game.c:97-130,165-267) on an SoA batch and read the tables back.
 */
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

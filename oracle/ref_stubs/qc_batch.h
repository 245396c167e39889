/* Stub written for oracle/_ref: re-declares ONLY the two batch structs that
 * old/chaos_game.c:174-193 dereferences (their real header is in the missing
 * bioinfo-libs submodule; field shapes from SURVEY.md Appendix A). */
#ifndef HPGQ_STUB_QC_BATCH_H
#define HPGQ_STUB_QC_BATCH_H
typedef struct fastq_batch {
  int num_reads;
  char *seq;
  char *quality;
  int *data_indices; /* num_reads + 1 prefix offsets */
} fastq_batch_t;
typedef struct status_batch {
  fastq_batch_t *read_p;
  int *read_status;
} status_batch_t;
#define VALID_READ 1
#define ONLY_VALID_READS 1
#define ALL_READS 0
#endif

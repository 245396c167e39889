/* stub for oracle/_ref */
#ifndef HPGQ_STUB_FASTQ_COMMONS_H
#define HPGQ_STUB_FASTQ_COMMONS_H
#define MAX_FULL_PATH_LENGTH 2048
#endif

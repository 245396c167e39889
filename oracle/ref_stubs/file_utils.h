/* stub for oracle/_ref: logging macros used by old/chaos_game.c */
#ifndef HPGQ_STUB_FILE_UTILS_H
#define HPGQ_STUB_FILE_UTILS_H
#include <stdio.h>
#include <stdlib.h>
#define LOG_FATAL(msg) do { fprintf(stderr, "%s", msg); exit(-1); } while (0)
#define LOG_DEBUG(msg) do { } while (0)
#define LOG_INFO(msg) do { } while (0)
#endif

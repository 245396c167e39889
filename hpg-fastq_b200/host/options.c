/*
 * options.c -- command-line surface of `hpg-fastq stats|filter|edit`.
 *
 * Same option names, defaults, validation messages and PARAMETERS CONFIGURATION display as the reference
 * (src/stats_options.c:60-290, src/filter_options.c, src/edit_options.c, src/commons_fastq.c:31-103);
 * argtable2 is replaced by a small long-option scanner (argtable2 is not on this box).
 */
#include "hpgq_host.h"

#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>

enum { A_LIT, A_INT, A_STR };

typedef struct hq_optdef {
  const char *shortn, *longn;
  int kind;
  const char *help;
  unsigned cmds; /* bit per hq_command */
  int ext;       /* 1: extension of this implementation */
} hq_optdef_t;

#define C_S 1u
#define C_F 2u
#define C_E 4u
#define C_ALL 7u

/* order = the reference's argtable order (src/stats_options.c:262-284) */
static const hq_optdef_t OPTS[] = {
    {"h", "help", A_LIT, "Help option", C_ALL, 0},
    {"f", "fastq-file", A_STR, "Input file name (FastQ format)", C_ALL, 0},
    {"o", "outdir", A_STR, "Output directory name", C_ALL, 0},
    {NULL, "num-threads", A_INT, "Number of threads", C_ALL, 0},
    {NULL, "batch-size", A_INT, "Batch size (in number of alignments)", C_ALL, 0},
    {NULL, "quality-encoding", A_STR, "Encoding for quality scores: phred33, phred64", C_S, 0},
    {NULL, "kmers", A_LIT, "Enable k-mers analysis (5-mer)", C_S, 0},
    {NULL, "read-length-range", A_STR, "Read length range, eg. 80,110", C_ALL, 0},
    {NULL, "read-quality-range", A_STR, "Read quality range, eg. 20,40", C_ALL, 0},
    {NULL, "left-length", A_INT, "Number of leftmost nucleotides to take into account to filter or trim", C_ALL, 0},
    {NULL, "left-quality-range", A_STR, "Quality range for the leftmost nucleotides, eg. 15,45", C_ALL, 0},
    {NULL, "right-length", A_INT, "Number of rightmost nucleotides to take into account to filter or trim", C_ALL, 0},
    {NULL, "right-quality-range", A_STR, "Quality range for the rightmost nucleotides, eg. 10,60", C_ALL, 0},
    {NULL, "max-N", A_INT, "Maximum number of Ns in the sequences", C_ALL, 0},
    {NULL, "max-out-of-quality", A_INT, "Maximum number of nucleotides out of the read quality range", C_ALL, 0},
    {NULL, "gpus", A_INT, "[B200] number of GPUs, batches are dealt round-robin (default 1)", C_ALL, 1},
    {NULL, "max-read-length", A_INT, "[B200] longest read to expect (skips the sizing pre-scan)", C_S, 1},
    {NULL, "kmers2", A_INT, "[B200] also count k-mers of this size (eg. 11)", C_S, 1},
    {NULL, "chaos-game", A_INT, "[B200] chaos-game frequency matrices of this word size (eg. 10)", C_S, 1},
    {NULL, "host-parse", A_LIT, "[B200] cut the FastQ records on the host instead of on the GPU", C_ALL, 1},
    {NULL, "chaos-game-signature", A_STR, "[B200] genomic signature file (.gs) for the chaos-game difference image", C_S, 1},
    /* paired-end input of the legacy generation (old/README:59-60, old/main_hpg_fastq_old.c:168-173): both mate
     * files in one invocation, "the outputs are calculated seperately for both files" (old/README:181) */
    {NULL, "fq1", A_STR, "[legacy] paired-end input, mate 1 (FastQ); with --fq2 instead of -f", C_ALL, 1},
    {NULL, "fq2", A_STR, "[legacy] paired-end input, mate 2 (FastQ)", C_ALL, 1},
    {NULL, "fastq1", A_STR, "[legacy] synonymous of --fq1", C_ALL, 1},
    {NULL, "fastq2", A_STR, "[legacy] synonymous of --fq2", C_ALL, 1},
    {NULL, "fq", A_STR, "[legacy] synonymous of --fastq-file", C_ALL, 1},
    {NULL, "fastq", A_STR, "[legacy] synonymous of --fastq-file", C_ALL, 1},
};
#define N_OPTS ((int)(sizeof OPTS / sizeof OPTS[0]))

static char *dupstr(const char *s) {
  size_t n = strlen(s) + 1;
  char *p = (char *)malloc(n);
  if (p) memcpy(p, s, n);
  return p;
}

static int path_exists(const char *p) {
  struct stat st;
  return p && stat(p, &st) == 0;
}

void hq_usage(const char *exec_name, const char *command_name, int command, FILE *out) {
  /* usage_argtable(), src/commons_fastq.c:22-27: syntax line then the glossary "%-50s\t%s\n" */
  fprintf(out, "Usage: %s %s\n", exec_name, command_name);
  for (int i = 0; i < N_OPTS; i++) {
    const hq_optdef_t *d = &OPTS[i];
    if (!(d->cmds & (1u << command))) continue;
    char left[96];
    const char *arg = d->kind == A_INT ? "=<int>" : d->kind == A_STR ? (i == 1 || i == 2 ? "=<file>" : "=<string>") : "";
    if (d->shortn) snprintf(left, sizeof left, "-%s, --%s%s", d->shortn, d->longn, arg);
    else snprintf(left, sizeof left, "--%s%s", d->longn, arg);
    fprintf(out, "  %-50s\t%s\n", left, d->help);
  }
}

static hq_options_t *options_new(const char *exec_name, const char *command_name, int command) {
  hq_options_t *o = (hq_options_t *)calloc(1, sizeof *o);
  if (!o) return NULL;
  o->command = command;
  o->num_threads = HQ_DEFAULT_NUM_THREADS;
  o->batch_size = HQ_DEFAULT_BATCH_SIZE;
  o->quality_encoding_value = 33;
  /* stats_options_new(), src/stats_options.c:27-52: every filter int starts as NO_VALUE */
  o->min_read_length = o->max_read_length = HPGQ_NO_VALUE;
  o->min_read_quality = o->max_read_quality = HPGQ_NO_VALUE;
  o->left_length = o->min_left_quality = o->max_left_quality = HPGQ_NO_VALUE;
  o->right_length = o->min_right_quality = o->max_right_quality = HPGQ_NO_VALUE;
  o->max_N = o->max_out_of_quality = HPGQ_NO_VALUE;
  o->exec_name = dupstr(exec_name);
  o->command_name = dupstr(command_name);
  o->gpus = 1;
  return o;
}

void hq_options_free(hq_options_t *o) {
  if (!o) return;
  free(o->quality_encoding_name);
  free(o->read_length_range);
  free(o->read_quality_range);
  free(o->left_quality_range);
  free(o->right_quality_range);
  free(o->cg_signature);
  free(o->in_filename);
  free(o->in_filename1);
  free(o->in_filename2);
  free(o->out_dirname);
  free(o->exec_name);
  free(o->command_name);
  free(o);
}

static int set_option(hq_options_t *o, int i, const char *val) {
  int iv = 0;
  if (OPTS[i].kind == A_INT) {
    char *end = NULL;
    long v = strtol(val, &end, 10);
    if (!val[0] || (end && *end)) {
      printf("\nError: invalid integer '%s' for --%s\n", val, OPTS[i].longn);
      return 0;
    }
    iv = (int)v;
  }
#define SETS(field) do { free(o->field); o->field = dupstr(val); } while (0)
  switch (i) {
    case 0: o->help = 1; break;
    case 1: SETS(in_filename); break;
    case 2: SETS(out_dirname); break;
    case 3: o->num_threads = iv; break;
    case 4: o->batch_size = iv; break;
    case 5: SETS(quality_encoding_name); break;
    case 6: o->kmers_on = 1; break;
    case 7: SETS(read_length_range); break;
    case 8: SETS(read_quality_range); break;
    case 9: o->left_length = iv; break;
    case 10: SETS(left_quality_range); break;
    case 11: o->right_length = iv; break;
    case 12: SETS(right_quality_range); break;
    case 13: o->max_N = iv; break;
    case 14: o->max_out_of_quality = iv; break;
    case 15: o->gpus = iv; break;
    case 16: o->max_read_len = iv; break;
    case 17: o->k2 = iv; break;
    case 18: o->cg_k = iv; break;
    case 19: o->host_parse = 1; break;
    case 20: SETS(cg_signature); break;
    case 21: case 23: SETS(in_filename1); break;
    case 22: case 24: SETS(in_filename2); break;
    case 25: case 26: SETS(in_filename); break;
    default: return 0;
  }
#undef SETS
  return 1;
}

hq_options_t *hq_options_parse(const char *exec_name, const char *command_name, int argc, char **argv) {
  int command = strcmp(command_name, "stats") == 0 ? HQ_CMD_STATS : strcmp(command_name, "filter") == 0 ? HQ_CMD_FILTER : HQ_CMD_EDIT;
  hq_options_t *o = options_new(exec_name, command_name, command);
  if (!o) return NULL;
  int bad = 0;
  for (int a = 1; a < argc && !bad; a++) {
    const char *s = argv[a];
    int found = -1;
    const char *val = NULL;
    if (s[0] == '-' && s[1] == '-') {
      const char *name = s + 2, *eq = strchr(name, '=');
      size_t nl = eq ? (size_t)(eq - name) : strlen(name);
      for (int i = 0; i < N_OPTS; i++)
        if ((OPTS[i].cmds & (1u << command)) && strlen(OPTS[i].longn) == nl && strncmp(OPTS[i].longn, name, nl) == 0) found = i;
      if (eq) val = eq + 1;
    } else if (s[0] == '-' && s[1] && !s[2]) {
      for (int i = 0; i < N_OPTS; i++)
        if ((OPTS[i].cmds & (1u << command)) && OPTS[i].shortn && OPTS[i].shortn[0] == s[1]) found = i;
    }
    if (found < 0) {
      printf("\nError: invalid option \"%s\"\n", s);
      bad = 1;
      break;
    }
    if (OPTS[found].kind != A_LIT && !val) {
      if (a + 1 >= argc) {
        printf("\nError: option \"%s\" requires an argument\n", s);
        bad = 1;
        break;
      }
      val = argv[++a];
    }
    if (!set_option(o, found, val ? val : "")) bad = 1;
  }
  if (bad || o->help) {
    hq_usage(exec_name, command_name, command, stdout);
    hq_options_free(o);
    return NULL;
  }
  return o;
}

int hq_parse_range(int *min, int *max, const char *range, const char *msg) {
  /* same accepted forms and messages as parse_range(), src/commons_fastq.c:31-103:
   * "a,b"  "a,"  ",b"  "a" ; negative or inverted ranges are errors */
  *min = HPGQ_NO_VALUE;
  *max = HPGQ_NO_VALUE;
  if (!range) return 1;
  int lmin = HPGQ_NO_VALUE, lmax = HPGQ_NO_VALUE;
  const char *comma = strchr(range, ',');
  char left[64];
  size_t nl = comma ? (size_t)(comma - range) : strlen(range);
  if (nl >= sizeof left) nl = sizeof left - 1;
  memcpy(left, range, nl);
  left[nl] = 0;
  if (comma && comma[1] && sscanf(comma + 1, "%d", &lmax) != 1) {
    printf("\nError: Invalid maximum value in the %s (%s)\n", msg, range);
    return 0;
  }
  if ((!comma || nl > 0) && sscanf(left, "%d", &lmin) != 1) {
    printf("\nError: Invalid minimum value in the %s (%s)\n", msg, range);
    return 0;
  }
  if (lmin != HPGQ_NO_VALUE && lmin < 0) {
    printf("\nError: Invalid %s (%s). Minimum value (%i) must be greater than 0\n", msg, range, lmin);
    return 0;
  }
  if (lmax != HPGQ_NO_VALUE && lmax < 0) {
    printf("\nError: Invalid %s (%s). Maximum value (%i) must be greater than 0\n", msg, range, lmax);
    return 0;
  }
  if (lmin != HPGQ_NO_VALUE && lmax != HPGQ_NO_VALUE && lmin > lmax) {
    printf("\nError: Invalid %s (%s). Maximum value (%i) must be greater than minimum value (%i)\n", msg, range, lmax, lmin);
    return 0;
  }
  *min = lmin;
  *max = lmax;
  return 1;
}

int hq_options_validate(hq_options_t *o) {
  /* stats_options_validate(), src/stats_options.c:110-160 (filter/edit siblings are the same minus the encoding) */
  if (o->in_filename1 || o->in_filename2) {
    /* paired end: both mates are mandatory (old/README:59-60), and they replace -f */
    if (!o->in_filename1 || !o->in_filename2 || o->in_filename) {
      printf("\nError: paired-end input needs both --fq1 and --fq2 (and no --fastq-file)\n");
      hq_usage(o->exec_name, o->command_name, o->command, stdout);
      return 0;
    }
    if (!path_exists(o->in_filename1) || !path_exists(o->in_filename2)) {
      printf("\nError: Input file name not found !\n");
      hq_usage(o->exec_name, o->command_name, o->command, stdout);
      return 0;
    }
    o->paired = 1;
    o->in_filename = dupstr(o->in_filename1);
  }
  if (!path_exists(o->in_filename)) {
    printf("\nError: Input file name not found !\n");
    hq_usage(o->exec_name, o->command_name, o->command, stdout);
    return 0;
  }
  if (!path_exists(o->out_dirname)) {
    free(o->out_dirname);
    o->out_dirname = dupstr(".");
  }
  if (o->quality_encoding_name) {
    if (strcmp(o->quality_encoding_name, HQ_PHRED33_NAME) == 0) {
      o->quality_encoding_value = 33;
    } else if (strcmp(o->quality_encoding_name, HQ_PHRED64_NAME) == 0) {
      o->quality_encoding_value = 64;
    } else {
      printf("\nError: Invalid quality encoding value (%s). Valid values: %s, %s\n", o->quality_encoding_name,
             HQ_PHRED33_NAME, HQ_PHRED64_NAME);
      hq_usage(o->exec_name, o->command_name, o->command, stdout);
      return 0;
    }
  } else {
    o->quality_encoding_name = dupstr(HQ_PHRED33_NAME);
    o->quality_encoding_value = 33;
  }
  if (!hq_parse_range(&o->min_read_length, &o->max_read_length, o->read_length_range, "read length range") ||
      !hq_parse_range(&o->min_read_quality, &o->max_read_quality, o->read_quality_range, "read quality range") ||
      !hq_parse_range(&o->min_left_quality, &o->max_left_quality, o->left_quality_range, "left quality range") ||
      !hq_parse_range(&o->min_right_quality, &o->max_right_quality, o->right_quality_range, "right quality range")) {
    hq_usage(o->exec_name, o->command_name, o->command, stdout);
    return 0;
  }
  if (o->gpus < 1 || o->gpus > 16 || o->batch_size < 1 || o->k2 < 0 || o->k2 > 13 || o->cg_k < 0 || o->cg_k > 12) {
    printf("\nError: --gpus must be 1..16, --batch-size >= 1, --kmers2 <= 13, --chaos-game <= 12\n");
    hq_usage(o->exec_name, o->command_name, o->command, stdout);
    return 0;
  }
  return 1;
}

int hq_options_display(hq_options_t *o, FILE *out) {
  /* stats_options_display() src/stats_options.c:164-226, filter_options_display() src/filter_options.c:142-196,
   * edit_options_display() src/edit_options.c:166-232 -- same text, same filter_on / edit_count rules */
  const int is_edit = o->command == HQ_CMD_EDIT;
  fprintf(out, "PARAMETERS CONFIGURATION\n");
  fprintf(out, "=================================================\n");
  fprintf(out, "Command name : %s\n", o->command_name);
  if (is_edit) fprintf(out, "\nMain edit_options\n");
  else fprintf(out, "\nMain options\n");
  fprintf(out, "\tFastQ input filename : %s\n", o->in_filename);
  fprintf(out, "\tOutput dirname       : %s\n", o->out_dirname);
  if (o->command == HQ_CMD_STATS) {
    fprintf(out, "\tK-mers (5-mer)       : %s\n", o->kmers_on ? "Enabled" : "Disabled");
    fprintf(out, "\tQuality encoding     : %s\n", o->quality_encoding_name);
  }
  int edit_count = 0, filter_count = 0;
  const int left_on = o->left_length != HPGQ_NO_VALUE && o->left_quality_range;
  const int right_on = o->right_length != HPGQ_NO_VALUE && o->right_quality_range;
  if (is_edit) {
    fprintf(out, "\nEdit options\n");
    if (left_on) {
      edit_count++;
      fprintf(out, "\tTrim left length         : %i nucleotides\n", o->left_length);
      fprintf(out, "\tTrim left quality range  : %s\n", o->left_quality_range);
    }
    if (right_on) {
      edit_count++;
      fprintf(out, "\tTrim right length        : %i nucleotides\n", o->right_length);
      fprintf(out, "\tTrim right quality range : %s\n", o->right_quality_range);
    }
    if (edit_count == 0) fprintf(out, "\tNone.\n\n");
  }
  fprintf(out, "\nFilter options\n");
  if (o->read_length_range) {
    filter_count++;
    fprintf(out, "\tRead length range   : %s\n", o->read_length_range);
  }
  if (o->read_quality_range) {
    filter_count++;
    fprintf(out, "\tRead quality range  : %s\n", o->read_quality_range);
  }
  if (!is_edit && left_on) {
    filter_count++;
    fprintf(out, "\tLeft length         : %i nucleotides\n", o->left_length);
    fprintf(out, "\tLeft quality range  : %s\n", o->left_quality_range);
  }
  if (!is_edit && right_on) {
    filter_count++;
    fprintf(out, "\tRight length        : %i nucleotides\n", o->right_length);
    fprintf(out, "\tRight quality range : %s\n", o->right_quality_range);
  }
  if (o->max_N != HPGQ_NO_VALUE) {
    filter_count++;
    fprintf(out, "\tMax. number of Ns   : %i\n", o->max_N);
  }
  if (o->max_out_of_quality != HPGQ_NO_VALUE && o->read_quality_range) {
    filter_count++;
    fprintf(out, "\tMax. out of quality : %i nucletotides\n", o->max_out_of_quality);
  }
  if (filter_count == 0) fprintf(out, is_edit ? "\tNone.\n\n" : "\tNone.\n");
  o->filter_on = filter_count != 0;
  o->edit_on = edit_count != 0;
  fprintf(out, "\nArchitecture options\n");
  fprintf(out, "\tNum. threads: %d\n", o->num_threads);
  fprintf(out, "\tBatch size  : %d alignments\n", o->batch_size);
  fprintf(out, "=================================================\n");
  if (o->command == HQ_CMD_FILTER && filter_count == 0) {
    fprintf(out, "\n\nNothing to filter, no filter options specified !\n\n");
    return 0;
  }
  if (is_edit && edit_count == 0) {
    fprintf(out, "\n\nNothing to edit, no edit options specified !\n\n");
    return 0;
  }
  return 1;
}

void hq_options_fill_unset(hq_options_t *o) {
#define UNSET(f, v) do { if (o->f == HPGQ_NO_VALUE) o->f = (v); } while (0)
  UNSET(min_read_length, HPGQ_MIN_VALUE);
  UNSET(max_read_length, HPGQ_MAX_VALUE);
  UNSET(min_read_quality, HPGQ_MIN_VALUE);
  UNSET(max_read_quality, HPGQ_MAX_VALUE);
  UNSET(max_out_of_quality, HPGQ_MAX_VALUE);
  UNSET(left_length, HPGQ_MIN_VALUE);
  UNSET(min_left_quality, HPGQ_MIN_VALUE);
  UNSET(max_left_quality, HPGQ_MAX_VALUE);
  UNSET(right_length, HPGQ_MIN_VALUE);
  UNSET(min_right_quality, HPGQ_MIN_VALUE);
  UNSET(max_right_quality, HPGQ_MAX_VALUE);
  UNSET(max_N, HPGQ_MAX_VALUE);
#undef UNSET
}

void hq_options_to_params(const hq_options_t *o, hpgq_params_t *p) {
  memset(p, 0, sizeof *p);
  /* the filter / edit tuples carry no encoding: Phred+33 there; `stats` forwards --quality-encoding */
  p->phred = o->command == HQ_CMD_STATS ? o->quality_encoding_value : 33;
  p->cg_base_quality = p->phred;
  hpgq_filter_opts_t *f = &p->filter;
  f->min_read_length = o->min_read_length;
  f->max_read_length = o->max_read_length;
  f->min_read_quality = o->min_read_quality;
  f->max_read_quality = o->max_read_quality;
  f->max_out_of_quality = o->max_out_of_quality;
  f->max_N = o->max_N;
  /* "unset" windows for a command that does not use them */
  f->left_length = HPGQ_MIN_VALUE;
  f->min_left_quality = HPGQ_MIN_VALUE;
  f->max_left_quality = HPGQ_MAX_VALUE;
  f->right_length = HPGQ_MIN_VALUE;
  f->min_right_quality = HPGQ_MIN_VALUE;
  f->max_right_quality = HPGQ_MAX_VALUE;
  p->edit.max_left_quality = p->edit.max_right_quality = HPGQ_MAX_VALUE;
  if (o->command == HQ_CMD_EDIT) {
    /* fastq_edit_options_new(9 ints) src/edit_fastq.c:148-151; the following filter has its window checks
     * disabled (src/edit_fastq.c:159-164) */
    p->work = HPGQ_WORK_EDIT | HPGQ_WORK_COMPACT | (o->filter_on ? HPGQ_WORK_FILTER : 0u);
    p->edit.left_length = o->left_length;
    p->edit.min_left_quality = o->min_left_quality;
    p->edit.max_left_quality = o->max_left_quality;
    p->edit.right_length = o->right_length;
    p->edit.min_right_quality = o->min_right_quality;
    p->edit.max_right_quality = o->max_right_quality;
  } else {
    /* fastq_filter_options_new(12 ints) src/stats_fastq.c:209-214, src/filter_fastq.c:140-145 */
    f->left_length = o->left_length;
    f->min_left_quality = o->min_left_quality;
    f->max_left_quality = o->max_left_quality;
    f->right_length = o->right_length;
    f->min_right_quality = o->min_right_quality;
    f->max_right_quality = o->max_right_quality;
    if (o->command == HQ_CMD_FILTER) {
      p->work = HPGQ_WORK_FILTER | HPGQ_WORK_COMPACT;
    } else {
      p->work = HPGQ_WORK_STATS | (o->filter_on ? HPGQ_WORK_FILTER : 0u) | (o->kmers_on ? HPGQ_WORK_KMERS : 0u) |
                (o->cg_k > 0 ? HPGQ_WORK_CHAOS : 0u);
      p->k2 = o->kmers_on ? o->k2 : 0;
      p->cg_k = o->cg_k;
    }
  }
}

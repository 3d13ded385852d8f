/* Shared command-line handling of the C drivers (tools/train_nn.c, tools/run_nn.c).
 * Flags as in the reference drivers (tests/train_nn.c:83-210): -h, -v (repeatable), -x,
 * -O n / -On (OpenMP threads), -B n (BLAS threads, accepted), -S n (CUDA streams, accepted);
 * first non-switch argument = configuration file, default ./nn.conf. */
#ifndef HPNN_TOOLS_NN_CLI_H
#define HPNN_TOOLS_NN_CLI_H
#include <libhpnn.h>

typedef struct {
  const char *conf;
  unsigned omp, blas, streams;
  int help, bad;
} nn_cli;

static unsigned cli_number(int argc, char **argv, int *i, const char *rest, char flag, nn_cli *c) {
  const char *v = rest[0] ? rest : (*i + 1 < argc ? argv[++(*i)] : "");
  while (*v == ' ' || *v == '\t') ++v;
  unsigned n = ISDIGIT(*v) ? (unsigned)strtoul(v, NULL, 10) : 0;
  if (n == 0) {
    _OUT(stderr, "syntax error: bad -%c parameter!\n", flag);
    c->bad = 1;
  }
  return n;
}

static nn_cli cli_parse(int argc, char **argv, int allow_dry) {
  nn_cli c = {NULL, 0, 0, 1, 0, 0};
  for (int i = 1; i < argc && !c.bad && !c.help; ++i) {
    const char *a = argv[i];
    if (a[0] != '-') {
      if (c.conf) c.bad = 1;
      else c.conf = a;
      continue;
    }
    for (int j = 1; a[j] && !c.bad; ++j) {
      const char f = a[j];
      if (f == 'h') {
        c.help = 1;
        break;
      } else if (f == 'v') {
        _NN(inc, verbose)();
      } else if (f == 'x' && allow_dry) {
        _NN(toggle, dry)();
      } else if (f == 'O' || f == 'B' || f == 'S') {
        unsigned n = cli_number(argc, argv, &i, a + j + 1, f, &c);
        if (f == 'O') c.omp = n;
        else if (f == 'B') c.blas = n;
        else c.streams = n;
        break; /* nothing may follow a numeric switch in the same word */
      } else {
        _OUT(stderr, "syntax error: unrecognized option!\n");
        c.bad = 1;
      }
    }
  }
  if (!c.conf) c.conf = "./nn.conf";
  return c;
}

static void cli_help(const char *prog, int train) {
  _OUT(stdout, "usage:  %s [-options] [input]\n", prog);
  _OUT(stdout, "options:\n-h \tdisplay this help;\n-v \tincrease verbosity;\n");
  if (train) _OUT(stdout, "-x \tdiscard results.\n");
  _OUT(stdout, "-O \tnumber of openMP threads.\n-B \tnumber of BLAS threads (accepted, unused).\n");
  _OUT(stdout, "-S \tnumber of CUDA streams (accepted; one persistent kernel per GPU).\n");
  _OUT(stdout, "input:  neural network .conf file (topology, kernel file or seed, directories).\n");
}

static void cli_apply(const nn_cli *c) {
  if (c->omp) _NN(set, omp_threads)(c->omp);
  if (c->blas) _NN(set, omp_blas)(c->blas);
  _NN(set, cuda_streams)(c->streams ? c->streams : 1);
}
#endif

/* train_nn: load nn.conf, write kernel.tmp, train one pass over the sample directory, write
 * kernel.opt.  Same contract as the reference driver (tests/train_nn.c:59-248), written against
 * the public libhpnn.h API only. */
#include <stdio.h>
#include <stdlib.h>

#include "nn_cli.h"

static int dump_to(nn_def *net, const char *path) {
  FILE *fp = fopen(path, "w");
  if (!fp) {
    _OUT(stderr, "FAILED to open %s for WRITE!\n", path);
    return -1;
  }
  _NN(dump, kernel)(net, fp);
  fclose(fp);
  return 0;
}

int main(int argc, char *argv[]) {
  int rc = -1;
  nn_def *net = NULL;
  _NN(init, all)(1);
  nn_cli cli = cli_parse(argc, argv, 1);
  if (cli.help || cli.bad) {
    cli_help("train_nn", 1);
    _NN(deinit, all)();
    return cli.bad ? -1 : 0;
  }
  cli_apply(&cli);
  net = _NN(load, conf)(cli.conf);
  if (!net) {
    _OUT(stderr, "FAILED to read NN configuration file! (ABORTING)\n");
  } else if (dump_to(net, "kernel.tmp") == 0) {
    if (!_NN(train, kernel)(net)) _OUT(stderr, "FAILED to train kernel!\n");
    else rc = dump_to(net, "kernel.opt");
  }
  if (net) {
    _NN(deinit, conf)(net);
    free(net);
  }
  _NN(deinit, all)();
  return rc;
}

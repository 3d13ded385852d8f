/* run_nn: load nn.conf (and the kernel it names), classify every file of the test directory.
 * Same contract as the reference driver (tests/run_nn.c:66-227). */
#include <stdio.h>
#include <stdlib.h>

#include "nn_cli.h"

int main(int argc, char *argv[]) {
  _NN(init, all)(1);
  nn_cli cli = cli_parse(argc, argv, 0);
  if (cli.help || cli.bad) {
    cli_help("run_nn", 0);
    _NN(deinit, all)();
    return cli.bad ? -1 : 0;
  }
  cli_apply(&cli);
  nn_def *net = _NN(load, conf)(cli.conf);
  if (!net) {
    _OUT(stderr, "FAILED to read NN configuration file! (ABORTING)\n");
    _NN(deinit, all)();
    return -1;
  }
  _NN(run, kernel)(net);
  _NN(deinit, conf)(net);
  free(net);
  _NN(deinit, all)();
  return 0;
}

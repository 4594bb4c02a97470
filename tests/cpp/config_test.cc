// CPU-only checks of the facade's host logic: linear::Config (text format of learn/base/arg_parser.h + defaults of
// learn/linear/config.proto) and the file pattern matching of learn/base/match_file.h. Prints one line per check.
#include <cstdio>
#include <string>

#include "linear/config.pb.h"

int main(int argc, char** argv) {
  dmlc::linear::Config c;
  std::string err;
  const std::string demo = "# demo conf\n\ntrain_data = \"learn/data/agaricus.txt.train\"\nval_data = \"learn/data/agaricus.txt.test\"\n"
                           "model_out = \"./agaricus_model\"\nmax_data_pass = 3\n";
  printf("parse %d\n", (int)c.ParseFromString(demo, &err));
  printf("train %s\n", c.train_data().c_str());
  printf("pass %d minibatch %d shuffle %d conc %d algo %d loss %d\n", c.max_data_pass(), c.minibatch(), c.rand_shuffle(),
         c.max_concurrency(), (int)c.algo(), (int)c.loss());
  printf("has_eta %d eta %g\n", (int)c.has_lr_eta(), c.lr_eta());
  printf("override %d\n", (int)c.ParseFromString("lr_eta=.1\nalgo=ADAGRAD\nminibatch: 500\n", &err));
  printf("has_eta %d eta %g algo %d minibatch %d\n", (int)c.has_lr_eta(), c.lr_eta(), (int)c.algo(), c.minibatch());
  printf("unknown %d\n", (int)c.ParseFromString("no_such_key = 1\n", &err));
  printf("badenum %d\n", (int)c.ParseFromString("algo = LBFGS\n", &err));
  return 0;
}

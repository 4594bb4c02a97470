// dmlc::linear::Progress: the six doubles a worker / server reports per minibatch and the row the
// scheduler prints from their sums (layout and format of learn/linear/progress.h:7-39).
#pragma once
#include <cstdio>
#include <string>
#include <vector>
namespace dmlc {
namespace linear {
struct Progress {
  Progress() : data(6) {}
  static std::string HeadStr() { return "  ttl #ex   inc #ex    |w|_0       logloss  accuracy     AUC"; }
  // accuracy and AUC are means of per-minibatch values, logloss a per-example mean (progress.h:21-23)
  std::string PrintStr() {
    ttl_ex += new_ex();
    nnz_w += new_w();
    if (new_ex() == 0) return "";
    char buf[256];
    snprintf(buf, sizeof(buf), "%8.3g  %8.3g  %11.6g  %8.6lf  %8.6lf  %8.6lf", ttl_ex, new_ex(), nnz_w,
             objv() / new_ex(), acc() / count(), auc() / count());
    return std::string(buf);
  }
  double& objv() { return data[0]; }
  double& acc() { return data[1]; }
  double& auc() { return data[2]; }
  double& count() { return data[3]; }
  double& new_ex() { return data[4]; }
  double& new_w() { return data[5]; }
  std::vector<double> data;
  double ttl_ex = 0, nnz_w = 0;
};
}  // namespace linear
}  // namespace dmlc

// dmlc::linear::Progress for the facade: the six numbers a worker / server reports per minibatch, the running totals
// the scheduler keeps, and the table row it prints (interface and column layout of learn/linear/progress.h:7-39;
// accuracy and AUC are means of per-minibatch values, logloss a per-example mean, :21-23).
#pragma once
#include <string>
#include <vector>

#include "wh_table.h"

namespace dmlc {
namespace linear {

struct Progress {
  enum Field { kObjv = 0, kAcc, kAuc, kCount, kNewEx, kNewW, kFields };
  std::vector<double> data = std::vector<double>(kFields, 0.0);
  double ttl_ex = 0, nnz_w = 0;   // totals over all rows printed so far

  double& objv() { return data[kObjv]; }
  double& acc() { return data[kAcc]; }
  double& auc() { return data[kAuc]; }
  double& count() { return data[kCount]; }
  double& new_ex() { return data[kNewEx]; }
  double& new_w() { return data[kNewW]; }

  static std::string HeadStr() {
    static const char* const kColumns[] = {"  ttl #ex", "   inc #ex", "    |w|_0", "       logloss", "  accuracy", "     AUC"};
    std::string head;
    for (const char* c : kColumns) head += c;
    return head;
  }

  /// folds `data` into the totals and renders the row; empty when no example was seen since the last row
  std::string PrintStr() {
    ttl_ex += data[kNewEx];
    nnz_w += data[kNewW];
    if (data[kNewEx] == 0) return std::string();
    const double n = data[kNewEx], mb = data[kCount];
    const wh::Cell row[] = {{"%8.3g", ttl_ex, "  "},          {"%8.3g", n, "  "},
                            {"%11.6g", nnz_w, "  "},          {"%8.6lf", data[kObjv] / n, "  "},
                            {"%8.6lf", data[kAcc] / mb, "  "}, {"%8.6lf", data[kAuc] / mb, ""}};
    return wh::FormatRow(row);
  }
};

}  // namespace linear
}  // namespace dmlc

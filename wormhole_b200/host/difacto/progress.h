// dmlc::difacto::Progress for the facade: eight numbers per minibatch, three running totals, one table row
// (interface and column layout of learn/difacto/progress.h:7-45).
#pragma once
#include <string>
#include <vector>

#include "wh_table.h"

namespace dmlc {
namespace difacto {

struct Progress {
  enum Field { kObjv = 0, kAuc, kObjvW, kCopc, kCount, kNewEx, kNewW, kNewV, kFields };
  std::vector<double> data = std::vector<double>(kFields, 0.0);
  double ttl_ex = 0, nnz_w = 0, nnz_V = 0;   // totals over all rows printed so far

  double& objv() { return data[kObjv]; }
  double& auc() { return data[kAuc]; }
  double& objv_w() { return data[kObjvW]; }
  double& copc() { return data[kCopc]; }
  double& count() { return data[kCount]; }
  double& new_ex() { return data[kNewEx]; }
  double& new_w() { return data[kNewW]; }
  double& new_V() { return data[kNewV]; }

  static std::string HeadStr() {
    static const char* const kColumns[] = {"  ttl #ex", "   inc #ex", " |  |w|_0", "  logloss_w", " |   |V|_0", "    logloss", "    AUC"};
    std::string head;
    for (const char* c : kColumns) head += c;
    return head;
  }

  /// folds `data` into the totals and renders the row; empty when no example was seen since the last row
  std::string PrintStr() {
    ttl_ex += data[kNewEx];
    nnz_w += data[kNewW];
    nnz_V += data[kNewV];
    if (data[kNewEx] == 0) return std::string();
    const double n = data[kNewEx];
    const wh::Cell row[] = {{"%9.4g", ttl_ex, "  "},  {"%7.2g", n, " | "},
                            {"%9.4g", nnz_w, "  "},   {"%6.4lf", data[kObjvW] / n, " | "},
                            {"%9.4g", nnz_V, "  "},   {"%7.5lf", data[kObjv] / n, "  "},
                            {"%7.5lf", data[kAuc] / data[kCount], " "}};
    return wh::FormatRow(row);
  }
};

}  // namespace difacto
}  // namespace dmlc

// Console-table helpers shared by the facade's Progress structs: a row is a list of printf-formatted cells joined by
// fixed separators, so that the column layouts of learn/linear/progress.h and learn/difacto/progress.h are data here.
#pragma once
#include <cstdio>
#include <string>

namespace wh {

struct Cell {
  const char* fmt;   // printf format of one double
  double value;
  const char* sep;   // text that follows the cell
};

template <size_t N>
inline std::string FormatRow(const Cell (&cells)[N]) {
  std::string row;
  char buf[64];
  for (const Cell& c : cells) {
    snprintf(buf, sizeof(buf), c.fmt, c.value);
    row += buf;
    row += c.sep;
  }
  return row;
}

}  // namespace wh

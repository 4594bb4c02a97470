// Stand-in for dmlc-core's src/data/strtonum.h (not vendored): the text parsers the reference itself ships
// (criteo_parser.h, adfea_parser.h) use libc atof / strtoull only. Test infrastructure only.
#pragma once
#include <stdlib.h>

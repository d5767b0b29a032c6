// Build shim for oracle/_ref: libear only needs boost::make_unique, which std provides.
#pragma once
#include <memory>
namespace boost { using std::make_unique; }

#pragma once
#include <string>

#include "tables.h"

namespace qmmx {
// Build the exact classification tables for one (level set, CONTACT_PROX). Returns 0 or -1 (+err text).
int build_tables(int L, const double* price, const uint8_t* solid, double prox, DevTables* T, std::string* err);
// float64 brute force of qmmx_monolithic.py:1543-1545: level index, or -1 when too far / no levels
int classify_bruteforce(int L, const double* price, double prox, float p);
// the table lookup exactly as the kernels perform it (host mirror, for tests)
int classify_tables(const DevTables* T, float p);
}  // namespace qmmx

#pragma once
#include <string>

#include "tables.h"

namespace qmmx {
// Build the exact classification tables for one (level set, CONTACT_PROX, Q_MIN_PROB). Returns 0 or -1 (+err).
int build_tables(int L, const double* price, const uint8_t* solid, double prox, double qmin, DevTables* T,
                 std::string* err);
// float64 brute force of qmmx_monolithic.py:1543-1545: level index, or -1 when too far / no levels
int classify_bruteforce(int L, const double* price, double prox, float p);
// the table lookup exactly as the kernels perform it (host mirror, for tests)
int classify_tables(const DevTables* T, float p);
// the class partition (tables.h): level touched (-1 far, -2 no class tables) and the levels whose contact window holds p,
// through the tables as the lean / tile kernels read them, and by float64 brute force (M:1543-1545, M:1572-1576)
int classify_tables_fine(const DevTables* T, float p, uint64_t* mask);
int classify_bruteforce_fine(int L, const double* price, double prox, float p, uint64_t* mask);
// compute_confidence (M:1415-1427) with direction known, plain IEEE float64
double confidence_f64(double dist, bool solid, int tc, double prox);
// host-evaluated sklearn-form LR gate: mask [n_lr][kMaxLevels], prob [n_lr][kMaxLevels][2][16] (or null)
void build_ml_tables(int L, const double* price, double sp, double qmin, const double* lr, int n_lr,
                     uint32_t* mask, double* prob);
// smallest float64 z with 1 / (1 + exp(-z)) >= target
double sigmoid_threshold(double target);
}  // namespace qmmx

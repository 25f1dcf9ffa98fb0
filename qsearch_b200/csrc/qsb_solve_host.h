// Host-side job description shared by the solver translation units (qsb_solve.cu is compiled three times, see build.py).
#pragma once
#include <vector>

#include "qsb_host.h"
#include "qsb_solver.cuh"

enum { QSB_SOLVE_LM = 0, QSB_SOLVE_LBFGS = 1 };

struct SolveJob {
  qsb_ctx* ctx; int n_programs; const qsb_program* const* progs; int n_instances; const int32_t* prog_of_instance;
  const double* x0; const double* target; qsb_solver_opts opts; int method;
  double* out_x; double* out_f; int32_t* out_iters; int32_t* out_status;
  bool target_unitary = false;  // checked once per call: the body-frame LM (qsb_k_lm3) needs it
};


// Levenberg-Marquardt on 5 qubits lives in its own translation unit (compile time)
int qsb_launch_solver_lm5(const SolveJob& job, qsb_batch* b, QsbSolveArgs& args, const QsbDevice& d);
int qsb_launch_solver_lm5t(const SolveJob& job, qsb_batch* b, QsbSolveArgs& args, const QsbDevice& d);  // TF32 Gram product
// The body-frame LM kernel as 4-warp CTAs, two per SM (n = 3, 4): for batches of more than one instance per SM.  Own unit too.
int qsb_launch_solver_lm_w4(const SolveJob& job, qsb_batch* b, QsbSolveArgs& args, const QsbDevice& d, int nq, bool tf32);

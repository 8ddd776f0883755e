// Argument block and launcher of the streaming weight-gradient / first-layer GEMMs of mlp_wide.cu, shared with the
// recurrent path (mlp_gru.cu), which uses the generic form for dW_ih and dW_hh.
#pragma once
#include "mlp_tc_common.cuh"

namespace mappo {
namespace tc {

struct WideArgs {
  NetLayout L;
  const float* params;
  const float* x; int ld; long long rows; const int64_t* index; int x_div;
  int vec;                       // 1: x rows are 16-byte aligned (float4 loads)
  int nhalf;                     // forward: output-column parts (1, 2 or 4 = grid.y): W1' is staged 64 / nhalf columns at a time
  float* z1;                     // [rows][64]   forward: out
  float* xstats;                 // [rows][2]    forward: out, weight gradient: in
  const float* dz1;              // [rows][64]   weight gradient: in
  float* grad_partials;          // weight gradient: partial rows [2 * nsplit][total]
  int nsplit, ncg, colw;         // weight gradient: row splits, column groups, columns per group (multiple of 16, <= 240)
  // generic form of the weight gradient  G[j][k] = sum_rows dz[row][j] x[row][k],  j < 64, k < D:
  int D, dz_ld, norm, total;     // x columns; dz row stride; 1: x rows are normalised with xstats; floats per partial row
  int out_w;                     // slot offset of G in a partial row
  int transposed, rot;           // 0: G stored [j][k] (row stride D); 1: stored [(k + rot) % D][j] (row stride 64)
  int out_b;                     // >= 0: slot offset of the column sums of x, stored at (k + rot) % D; -1: none
  // dual form (the recurrent chain's two weight gradients from ONE pass over the gradient rows): a second dz operand and a
  // second output; output 1 takes the x columns [0, win), output 2 the columns [off2, off2 + win), each stored at
  // (k' + rot) % win with k' counted from its window's first column.  win == 0: single form (window = D).
  const float* dz2; int win, off2, rot2, out_w2, out_b2;
};

int launch_wgrad_generic(WideArgs& w, cudaStream_t stream, const char* who);
bool wide_fits(int in_dim);      // the staged first layer (fc1_wide_kernel) has room for this many input columns

}  // namespace tc
}  // namespace mappo

// Flat parameter layout of one actor/critic network (see include/mappo_b200.h section 3) and the
// "folded" form the kernels compute with.
//
// Folding: every LayerNorm in the reference is  a = g (.) n + be  with n the normalised activations, and is
// always followed by a Linear:  z = W a + b = (W diag(g)) n + (b + W be) = W' n + b'.
// The kernels keep only n in shared memory and use W', b'.  Their gradient buffers hold, in the SLOT of W,
// G[j][k] = sum_r dz[r][j] n[r][k] and, in the slot of b, db[j] = sum_r dz[r][j]; the true gradients follow by
// the chain rule in the optimiser kernel (optim.cu):
//     dW[j][k] = g[k] G[j][k] + be[k] db[j],   dg[k] = sum_j W[j][k] G[j][k],   dbe[k] = sum_j W[j][k] db[j].
#pragma once
#include "common.cuh"

namespace mappo {

constexpr int kH = 64;          // hidden size this build supports
constexpr int kMaxIn = 64;      // widest input handled by the single-chunk first layer (fp32 FMA kernels)
constexpr int kWideFrom = 64;   // tensor-core path: from this width on the first layer runs outside the fused kernel (mlp_wide.cu)

constexpr int kMaxRecurrentN = 2;   // mappo_net_config.recurrent carries recurrent_N (0 = feed-forward)
constexpr int kMaxLayerN = 3;   // mlp.py:20-28: fc2 = layer_N clones of fc_h (hidden -> hidden blocks after fc1)

struct NetLayout {
  int in_dim, H, n_heads, head_dim[MAPPO_MAX_HEADS], a_total;
  int use_fn, use_relu, recurrent, is_critic;
  int layer_N;
  int wx[kMaxLayerN - 1], bx[kMaxLayerN - 1], gx[kMaxLayerN - 1], bex[kMaxLayerN - 1];   // fc2[1 ..] (layer_N > 1)
  // offsets into the flat buffer
  int fn_g, fn_b;               // [in] each (valid if use_fn)
  int w1, b1, g1, be1;          // fc1
  int w2, b2, g2, be2;          // fc2[0]
  int wih, whh, bih, bhh, gr, ber;   // rnn layer 0 and its LayerNorm (valid if recurrent)
  int wih1, whh1, bih1, bhh1;        // rnn layer 1 (recurrent_N = 2: nn.GRU(num_layers = 2), rnn.py:13); -1 otherwise
  int rnn_layers;                    // recurrent_N (0 if not recurrent)
  int wh[MAPPO_MAX_HEADS], bh[MAPPO_MAX_HEADS];
  int total;
};

__host__ __device__ inline NetLayout make_layout(const mappo_net_config& c) {
  NetLayout L;
  L.in_dim = c.in_dim; L.H = c.hidden; L.n_heads = c.n_heads; L.use_fn = c.use_feature_norm;
  L.use_relu = c.use_relu; L.recurrent = c.recurrent; L.is_critic = c.is_critic;
  int o = 0;
  const int H = c.hidden, D = c.in_dim;
  L.fn_g = L.fn_b = -1;
  if (c.use_feature_norm) { L.fn_g = o; o += D; L.fn_b = o; o += D; }
  L.w1 = o; o += H * D; L.b1 = o; o += H; L.g1 = o; o += H; L.be1 = o; o += H;
  L.w2 = o; o += H * H; L.b2 = o; o += H; L.g2 = o; o += H; L.be2 = o; o += H;
  L.layer_N = c.layer_N < 1 ? 1 : (c.layer_N > kMaxLayerN ? kMaxLayerN : c.layer_N);
  for (int l = 0; l < kMaxLayerN - 1; ++l) {
    L.wx[l] = L.bx[l] = L.gx[l] = L.bex[l] = -1;
    if (l + 1 < L.layer_N) { L.wx[l] = o; o += H * H; L.bx[l] = o; o += H; L.gx[l] = o; o += H; L.bex[l] = o; o += H; }
  }
  L.wih = L.whh = L.bih = L.bhh = L.gr = L.ber = -1;
  L.wih1 = L.whh1 = L.bih1 = L.bhh1 = -1;
  L.rnn_layers = c.recurrent < 0 ? 0 : (c.recurrent > kMaxRecurrentN ? kMaxRecurrentN : c.recurrent);
  if (c.recurrent) {     // nn.GRU.named_parameters(): weight_ih_l0, weight_hh_l0, bias_ih_l0, bias_hh_l0, then layer 1, then rnn.norm
    L.wih = o; o += 3 * H * H; L.whh = o; o += 3 * H * H; L.bih = o; o += 3 * H; L.bhh = o; o += 3 * H;
    if (L.rnn_layers > 1) {
      L.wih1 = o; o += 3 * H * H; L.whh1 = o; o += 3 * H * H; L.bih1 = o; o += 3 * H; L.bhh1 = o; o += 3 * H;
    }
    L.gr = o; o += H; L.ber = o; o += H;
  }
  L.a_total = 0;
  for (int h = 0; h < MAPPO_MAX_HEADS; ++h) {
    L.head_dim[h] = h < c.n_heads ? c.head_dim[h] : 0;
    L.wh[h] = L.bh[h] = -1;
    if (h < c.n_heads) {
      L.wh[h] = o; o += c.head_dim[h] * H; L.bh[h] = o; o += c.head_dim[h];
      L.a_total += c.head_dim[h];
    }
  }
  L.total = o;
  return L;
}

// offsets of hidden block l (0 = fc2[0]) and of the LayerNorm affine that precedes the heads
__host__ __device__ inline int layer_w(const NetLayout& L, int l) { return l == 0 ? L.w2 : L.wx[l - 1]; }
__host__ __device__ inline int layer_b(const NetLayout& L, int l) { return l == 0 ? L.b2 : L.bx[l - 1]; }
__host__ __device__ inline int layer_g(const NetLayout& L, int l) { return l == 0 ? L.g2 : L.gx[l - 1]; }
__host__ __device__ inline int layer_be(const NetLayout& L, int l) { return l == 0 ? L.be2 : L.bex[l - 1]; }

static inline int check_net(const mappo_net_config* c, const char* who) {
  MAPPO_REQUIRE(c, "%s: null net config", who);
  MAPPO_REQUIRE(c->in_dim > 0 && c->n_heads >= 1 && c->n_heads <= MAPPO_MAX_HEADS, "%s: bad in_dim/n_heads", who);
  if (c->hidden != kH) MAPPO_UNSUPPORTED("%s: hidden_size %d (this build: 64)", who, c->hidden);
  if (c->layer_N < 1 || c->layer_N > kMaxLayerN) MAPPO_UNSUPPORTED("%s: layer_N %d (this build: 1 .. %d)", who, c->layer_N, kMaxLayerN);
  if (c->recurrent < 0 || c->recurrent > kMaxRecurrentN) MAPPO_UNSUPPORTED("%s: recurrent_N %d (this build: 1 .. %d)", who, c->recurrent, kMaxRecurrentN);
  if (c->layer_N != 1 && c->recurrent) MAPPO_UNSUPPORTED("%s: layer_N %d with a recurrent network (this build: 1)", who, c->layer_N);
  if (c->layer_N != 1 && c->in_dim > kMaxIn) MAPPO_UNSUPPORTED("%s: layer_N %d with in_dim %d > %d", who, c->layer_N, c->in_dim, kMaxIn);
  int a = 0;
  for (int h = 0; h < c->n_heads; ++h) { MAPPO_REQUIRE(c->head_dim[h] >= 1, "%s: head_dim", who); a += c->head_dim[h]; }
  MAPPO_REQUIRE(a <= 16, "%s: at most 16 action logits in total", who);
  MAPPO_REQUIRE(!c->is_critic || (c->n_heads == 1 && c->head_dim[0] == 1), "%s: critic has one scalar head", who);
  return 0;
}

}  // namespace mappo

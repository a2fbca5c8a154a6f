// PokerGPU on the reference-shaped tensors (include/pulse_envs.h: PulsePokerState).  One fused
// launch per step; the reference's internal methods are the same kernel with one phase bit, so the
// white-box tests of the reference (tests/poker/*) can drive every sub-step.
#include "poker_compat_body.cuh"

namespace {

template <int MAXP>
__global__ void __launch_bounds__(128)
poker_compat_step_kernel(PulsePokerState S, PokerParams pr, const int64_t* __restrict__ actions,
                         const int32_t* __restrict__ actor_idx, uint32_t phases, float* __restrict__ obs_out,
                         float* __restrict__ reward_out, uint8_t* __restrict__ done_out, int n) {
  for (int g = blockIdx.x * blockDim.x + threadIdx.x; g < n; g += gridDim.x * blockDim.x)
    compat_step_one<MAXP>(S, pr, actions, actor_idx, phases, obs_out, reward_out, done_out, g);
}

template <int MAXP>
__global__ void __launch_bounds__(128)
poker_compat_reset_kernel(PulsePokerState S, PokerParams pr, const int32_t* __restrict__ prefixed,
                          uint64_t seed, uint64_t env_base, uint32_t hand, int rotation, int first_reset,
                          int starting_bbs, int max_bbs, float* __restrict__ obs_out, int n) {
  for (int g = blockIdx.x * blockDim.x + threadIdx.x; g < n; g += gridDim.x * blockDim.x)
    compat_reset_one<MAXP>(S, pr, prefixed, seed, env_base, hand, rotation, first_reset, starting_bbs, max_bbs,
                           obs_out, g);
}

__global__ void poker_eval_batch_kernel(const int32_t* __restrict__ cards, int nc, int32_t* __restrict__ out, int n) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    uint32_t m[4] = {0, 0, 0, 0};
    for (int c = 0; c < nc; ++c) pe_add_card(cards[(size_t)i * nc + c], m);
    out[i] = pe_eval_suits(m[0], m[1], m[2], m[3]);
  }
}

int check_cfg(const PulsePokerState* st, const PulsePokerConfig* cfg) {
  if (!st || !cfg) { pulse_set_error("null state/config"); return PULSE_EINVAL; }
  if (cfg->n_players < 2 || cfg->n_players > 10) { pulse_set_error("n_players must be in 2..10"); return PULSE_EINVAL; }
  if (cfg->active_players < 2 || cfg->active_players > cfg->n_players) { pulse_set_error("active_players must be in 2..n_players"); return PULSE_EINVAL; }
  if (cfg->max_players < cfg->active_players || cfg->max_players > 16) { pulse_set_error("max_players must be in active_players..16"); return PULSE_EINVAL; }
  if (cfg->n_games < 0) { pulse_set_error("n_games < 0"); return PULSE_EINVAL; }
  return PULSE_OK;
}

PokerParams make_params(const PulsePokerConfig* c) {
  PokerParams p; p.w1 = c->w1; p.w2 = c->w2; p.K = c->K; p.alpha = c->alpha;
  p.P = c->n_players; p.AP = c->active_players; p.M = c->max_players;
  return p;
}

}  // namespace

#define PK_DISPATCH(P, CALL)                  \
  do {                                        \
    if ((P) <= 2) { constexpr int MP = 2; CALL; }        \
    else if ((P) <= 4) { constexpr int MP = 4; CALL; }   \
    else if ((P) <= 6) { constexpr int MP = 6; CALL; }   \
    else { constexpr int MP = 10; CALL; }                \
  } while (0)

extern "C" {

int pulse_poker_reset(const PulsePokerState* st, const PulsePokerConfig* cfg, const int32_t* prefixed,
                      uint64_t seed, uint64_t env_base, uint32_t hand_counter, int rotation, int first_reset,
                      float* obs_out, cudaStream_t stream) {
  int rc = check_cfg(st, cfg);
  if (rc) return rc;
  if (cfg->n_games == 0) return PULSE_OK;
  PokerParams pr = make_params(cfg);
  const int block = 128, grid = pulse_grid_for(cfg->n_games, block);
  PK_DISPATCH(pr.P, (poker_compat_reset_kernel<MP><<<grid, block, 0, stream>>>(
                        *st, pr, prefixed, seed, env_base, hand_counter, rotation, first_reset,
                        cfg->starting_bbs, cfg->max_bbs, obs_out, cfg->n_games)));
  return pulse_check_launch("pulse_poker_reset");
}

int pulse_poker_step(const PulsePokerState* st, const PulsePokerConfig* cfg, const int64_t* actions,
                     const int32_t* actor_idx, uint32_t phases, float* obs_out, float* reward_out,
                     uint8_t* done_out, cudaStream_t stream) {
  int rc = check_cfg(st, cfg);
  if (rc) return rc;
  if (cfg->n_games == 0) return PULSE_OK;
  PULSE_REQUIRE(!(phases & (PK_PH_EXECUTE | PK_PH_REWARD)) || actions, "actions is null");
  PULSE_REQUIRE(!(phases & PK_PH_REWARD) || reward_out, "reward_out is null");
  PULSE_REQUIRE(!(phases & PK_PH_OBS) || obs_out, "obs_out is null");
  PokerParams pr = make_params(cfg);
  const int block = 128, grid = pulse_grid_for(cfg->n_games, block);
  PK_DISPATCH(pr.P, (poker_compat_step_kernel<MP><<<grid, block, 0, stream>>>(
                        *st, pr, actions, actor_idx, phases, obs_out, reward_out, done_out, cfg->n_games)));
  return pulse_check_launch("pulse_poker_step");
}

int pulse_poker_eval_batch(const int32_t* cards, int n_cards, int32_t* out, int n, cudaStream_t stream) {
  PULSE_REQUIRE(n_cards >= 5 && n_cards <= 7, "n_cards must be 5..7");
  PULSE_REQUIRE(n >= 0, "n < 0");
  if (n == 0) return PULSE_OK;
  PULSE_REQUIRE(cards && out, "null buffer");
  poker_eval_batch_kernel<<<pulse_grid_for(n, 256), 256, 0, stream>>>(cards, n_cards, out, n);
  return pulse_check_launch("pulse_poker_eval_batch");
}
}

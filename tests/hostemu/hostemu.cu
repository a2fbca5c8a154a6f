// TEST INFRASTRUCTURE ONLY.  Compiles the per-env bodies of the CUDA kernels (the __host__
// __device__ functions in pulselib_b200/csrc/*_body.cuh) for the CPU and loops them over the batch,
// so the CPU test suite can check the kernel logic against the oracle without a GPU.  Not part of
// libpulse_envs.so and never used by the product.
#include "poker_compat_body.cuh"
#include "poker_packed_body.cuh"

void pulse_set_error(const char*, ...) {}
int pulse_check_launch(const char*) { return 0; }

static PokerParams mk(const PulsePokerConfig* c) {
  PokerParams p; p.w1 = c->w1; p.w2 = c->w2; p.K = c->K; p.alpha = c->alpha;
  p.P = c->n_players; p.AP = c->active_players; p.M = c->max_players;
  return p;
}

#define EMU_DISPATCH(P, CALL)                            \
  do {                                                   \
    if ((P) <= 2) { constexpr int MP = 2; CALL; }        \
    else if ((P) <= 4) { constexpr int MP = 4; CALL; }   \
    else if ((P) <= 6) { constexpr int MP = 6; CALL; }   \
    else { constexpr int MP = 10; CALL; }                \
  } while (0)

extern "C" {
int emu_poker_reset(const PulsePokerState* st, const PulsePokerConfig* cfg, const int32_t* prefixed, uint64_t seed,
                    uint64_t env_base, uint32_t hand, int rotation, int first_reset, float* obs_out) {
  PokerParams pr = mk(cfg);
  for (int g = 0; g < cfg->n_games; ++g)
    EMU_DISPATCH(pr.P, (compat_reset_one<MP>(*st, pr, prefixed, seed, env_base, hand, rotation, first_reset,
                                             cfg->starting_bbs, cfg->max_bbs, obs_out, g)));
  return 0;
}
int emu_poker_step(const PulsePokerState* st, const PulsePokerConfig* cfg, const int64_t* actions,
                   const int32_t* actor_idx, uint32_t phases, float* obs_out, float* reward_out, uint8_t* done_out) {
  PokerParams pr = mk(cfg);
  for (int g = 0; g < cfg->n_games; ++g)
    EMU_DISPATCH(pr.P, (compat_step_one<MP>(*st, pr, actions, actor_idx, phases, obs_out, reward_out, done_out, g)));
  return 0;
}

int emu_poker_packed_reset(const PulsePokerPacked* st, const PulsePokerConfig* cfg, const int32_t* prefixed,
                           uint64_t seed, uint64_t env_base, uint32_t hand, int rotation, int first_reset,
                           float* obs_out) {
  PokerParams pr = mk(cfg);
  const int W = 13 + 3 * (pr.M - 1);
  for (long long g = 0; g < cfg->n_games; ++g)
    EMU_DISPATCH(pr.P, (packed_reset_one<MP>(*st, pr, prefixed, seed, env_base, hand, rotation, first_reset,
                                             cfg->starting_bbs, cfg->max_bbs, obs_out ? obs_out + g * W : nullptr, g,
                                             (long long)cfg->n_games)));
  return 0;
}
int emu_poker_packed_step(const PulsePokerPacked* st, const PulsePokerConfig* cfg, const int64_t* actions,
                          float* obs_out, float* reward_out, uint8_t* done_out, int auto_reset, uint64_t seed,
                          uint64_t env_base) {
  PokerParams pr = mk(cfg);
  const int W = 13 + 3 * (pr.M - 1);
  for (long long g = 0; g < cfg->n_games; ++g)
    EMU_DISPATCH(pr.P, (packed_step_one<MP>(*st, pr, actions, obs_out + g * W, reward_out, done_out, auto_reset, seed,
                                            env_base, cfg->starting_bbs, cfg->max_bbs, g, (long long)cfg->n_games)));
  return 0;
}
int emu_poker_unpack(const PulsePokerPacked* st, const PulsePokerState* compat, const PulsePokerConfig* cfg) {
  PokerParams pr = mk(cfg);
  for (long long g = 0; g < cfg->n_games; ++g)
    EMU_DISPATCH(pr.P, (packed_unpack_one<MP>(*st, *compat, pr, g, (long long)cfg->n_games)));
  return 0;
}
}

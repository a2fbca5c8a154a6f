// Blackjack kernels and their C ABI (include/pulse_envs.h).  Per-env logic and the HBM layout are in
// blackjack_body.cuh.  One thread per env, one launch per step; the dealer's draw loop
// (blackjack.py:141-160, a host-synchronising `while .any()` in the reference) runs inside the thread.
#include "blackjack_body.cuh"

namespace {
constexpr int kBlock = 256;

__global__ void __launch_bounds__(kBlock)
blackjack_reset_kernel(uint32_t* __restrict__ st, uint8_t* __restrict__ deck, const uint8_t* __restrict__ injected,
                       uint64_t seed, uint64_t env_base, uint32_t episode, int32_t* __restrict__ obs,
                       int32_t* __restrict__ rewards, uint8_t* __restrict__ terminated, long long n) {
  for (long long g = (long long)blockIdx.x * kBlock + threadIdx.x; g < n; g += (long long)gridDim.x * kBlock)
    bj_reset_one(st, deck, injected, seed, env_base, episode, obs, rewards, terminated, g, n);
}

__global__ void __launch_bounds__(kBlock)
blackjack_step_kernel(uint32_t* __restrict__ st, const uint8_t* __restrict__ deck, const int64_t* __restrict__ actions,
                      int32_t* __restrict__ obs, int32_t* __restrict__ rewards, uint8_t* __restrict__ terminated,
                      long long n) {
  for (long long g = (long long)blockIdx.x * kBlock + threadIdx.x; g < n; g += (long long)gridDim.x * kBlock)
    bj_step_one(st, deck, actions, obs, rewards, terminated, g, n);
}

// K steps per launch (SURVEY.md 8(b) "K steps per launch"): at the reference's own batch size (4096 envs, config 1)
// a step is ~0.2 us of device work behind ~3 us of launch, so the loop over steps moves inside the kernel.  The env's
// state word stays in registers between steps; actions are [K, N], outputs [K, N, ...] (step-major, so every store of
// a warp is contiguous).  reset_first: deal a fresh hand (BlackJack.reset) before the first step.
template <class ActT>
__global__ void __launch_bounds__(128)
blackjack_rollout_kernel(uint32_t* __restrict__ st, uint8_t* __restrict__ deck, const ActT* __restrict__ actions, int K,
                         int n_hands, int reset_first, uint64_t seed, uint64_t env_base, uint32_t episode,
                         int32_t* __restrict__ reset_obs, int32_t* __restrict__ obs, int32_t* __restrict__ rewards,
                         uint8_t* __restrict__ terminated, long long n) {
  // n_hands hands of K steps each; hand h > 0 (and hand 0 when reset_first) starts with a fresh deal from the tape at
  // episode + (number of deals so far); actions / outputs are [n_hands * K, N, ...], reset_obs [n_hands, N, 3]
  for (long long g = (long long)blockIdx.x * 128 + threadIdx.x; g < n; g += (long long)gridDim.x * 128) {
    BjState s;
    uint32_t ep = episode;
    for (int h = 0; h < n_hands; ++h) {
      if (h > 0 || reset_first) {
        s = bj_reset_state(st, deck, nullptr, seed, env_base, ep++, g, n);
        if (reset_obs) {
          int32_t* ro = reset_obs + 3 * ((long long)h * n + g);
          ro[0] = s.psum; ro[1] = s.ace; ro[2] = s.upcard;
        }
      } else {
        s = bj_unpack_word(st[g]);
      }
      const long long k0 = (long long)h * K;
      long long a = K > 0 ? (long long)actions[k0 * n + g] : 0;
      for (int k = 0; k < K; ++k) {
        const long long a_next = k + 1 < K ? (long long)actions[(k0 + k + 1) * n + g] : 0;   // one step ahead
        const int reward = bj_step_state(s, a, deck, g, n);
        const long long o = (k0 + k) * n + g;
        obs[3 * o] = s.psum; obs[3 * o + 1] = s.ace; obs[3 * o + 2] = s.upcard;
        rewards[o] = reward; terminated[o] = s.term;
        a = a_next;
      }
    }
    st[g] = bj_pack(s);
  }
}

__global__ void __launch_bounds__(kBlock)
blackjack_unpack_kernel(const uint32_t* __restrict__ st, const uint8_t* __restrict__ deck, BjCompat C, long long n) {
  for (long long g = (long long)blockIdx.x * kBlock + threadIdx.x; g < n; g += (long long)gridDim.x * kBlock)
    bj_unpack_one(st, deck, C, g, n);
}
}  // namespace

extern "C" {

int pulse_blackjack_reset(uint32_t* st, uint8_t* deck, const uint8_t* injected_decks_or_null, uint64_t seed,
                          uint64_t env_base, uint32_t episode, int32_t* obs_out, int32_t* rewards_out,
                          uint8_t* terminated_out, int n_envs, cudaStream_t stream) {
  PULSE_REQUIRE(n_envs >= 0, "n_envs < 0");
  if (n_envs == 0) return PULSE_OK;
  PULSE_REQUIRE(st && deck && obs_out && rewards_out && terminated_out, "null buffer");
  blackjack_reset_kernel<<<pulse_grid_for(n_envs, kBlock), kBlock, 0, stream>>>(
      st, deck, injected_decks_or_null, seed, env_base, episode, obs_out, rewards_out, terminated_out, n_envs);
  return pulse_check_launch("pulse_blackjack_reset");
}

int pulse_blackjack_step(uint32_t* st, const uint8_t* deck, const int64_t* actions, int32_t* obs_out,
                         int32_t* rewards_out, uint8_t* terminated_out, int n_envs, cudaStream_t stream) {
  PULSE_REQUIRE(n_envs >= 0, "n_envs < 0");
  if (n_envs == 0) return PULSE_OK;
  PULSE_REQUIRE(st && deck && actions && obs_out && rewards_out && terminated_out, "null buffer");
  blackjack_step_kernel<<<pulse_grid_for(n_envs, kBlock), kBlock, 0, stream>>>(st, deck, actions, obs_out, rewards_out,
                                                                               terminated_out, n_envs);
  return pulse_check_launch("pulse_blackjack_step");
}

int pulse_blackjack_rollout(uint32_t* st, uint8_t* deck, const void* actions, int action_bytes, int n_steps, int n_hands,
                            int reset_first, uint64_t seed, uint64_t env_base, uint32_t episode,
                            int32_t* reset_obs_out_or_null, int32_t* obs_out, int32_t* rewards_out,
                            uint8_t* terminated_out, int n_envs, cudaStream_t stream) {
  PULSE_REQUIRE(n_envs >= 0 && n_steps >= 0 && n_hands >= 1, "n_envs / n_steps < 0 or n_hands < 1");
  PULSE_REQUIRE(action_bytes == 1 || action_bytes == 8, "actions must be uint8 or int64");
  if (n_envs == 0 || (n_steps == 0 && !reset_first && n_hands == 1)) return PULSE_OK;
  PULSE_REQUIRE(st && deck && (actions || n_steps == 0) && obs_out && rewards_out && terminated_out, "null buffer");
  const int grid = pulse_grid_for(n_envs, 128);
  if (action_bytes == 1)
    blackjack_rollout_kernel<uint8_t><<<grid, 128, 0, stream>>>(st, deck, (const uint8_t*)actions, n_steps, n_hands, reset_first, seed,
                                                                 env_base, episode, reset_obs_out_or_null, obs_out, rewards_out,
                                                                 terminated_out, n_envs);
  else
    blackjack_rollout_kernel<int64_t><<<grid, 128, 0, stream>>>(st, deck, (const int64_t*)actions, n_steps, n_hands, reset_first, seed,
                                                                 env_base, episode, reset_obs_out_or_null, obs_out, rewards_out,
                                                                 terminated_out, n_envs);
  return pulse_check_launch("pulse_blackjack_rollout");
}

int pulse_blackjack_unpack(const uint32_t* st, const uint8_t* deck, const PulseBlackjackCompat* out, int n_envs,
                           cudaStream_t stream) {
  PULSE_REQUIRE(n_envs >= 0, "n_envs < 0");
  if (n_envs == 0) return PULSE_OK;
  PULSE_REQUIRE(st && deck && out, "null buffer");
  BjCompat C;
  C.decks = out->decks; C.deck_positions = out->deck_positions; C.players_cards = out->players_cards;
  C.players_card_idx = out->players_card_idx; C.player_card_sums = out->player_card_sums; C.has_ace = out->has_ace;
  C.dealer_cards = out->dealer_cards; C.dealer_card_idx = out->dealer_card_idx; C.dealer_upcard = out->dealer_upcard;
  C.dealer_card_sums = out->dealer_card_sums; C.dealer_has_ace = out->dealer_has_ace;
  blackjack_unpack_kernel<<<pulse_grid_for(n_envs, kBlock), kBlock, 0, stream>>>(st, deck, C, n_envs);
  return pulse_check_launch("pulse_blackjack_unpack");
}
}

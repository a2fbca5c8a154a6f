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

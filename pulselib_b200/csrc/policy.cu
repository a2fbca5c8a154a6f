// build_actions + the scripted GPU opponents as one launch over the observation tensor
// (rules and reference lines: poker_policy.cuh).  One thread per table: seat_idx (4 B,
// coalesced) picks the seat's agent kind from a 16-byte kernel argument; scripted kinds read the acting
// seat's hole cards (obs columns 5..6; column 9 only for SmallBall) and write the int64 action; tables
// whose seat is PULSE_AGENT_EXTERNAL keep the action the caller put there (the Q-network's).
// The packed engine does the same inside its step kernel from registers (poker_packed.cu) -- this entry is
// for callers that hold the reference-shaped observation, like the reference's own loop
// (scripts/Poker/trainGPU.py:79-83).
#include "poker_policy.cuh"

namespace {
constexpr int kBlock = 256;

__global__ void __launch_bounds__(kBlock)
poker_policy_kernel(const float* __restrict__ obs, int W, const int32_t* __restrict__ seat_idx, PulseSeatKinds kinds,
                    const uint32_t* __restrict__ chance2, uint64_t seed, uint64_t env_base, uint32_t counter,
                    int64_t* __restrict__ actions, long long n) {
  for (long long g = (long long)blockIdx.x * kBlock + threadIdx.x; g < n; g += (long long)gridDim.x * kBlock) {
    const int seat = seat_idx[g];
    const int kind = (seat >= 0 && seat < 16) ? kinds.kind[seat] : PULSE_AGENT_EXTERNAL;
    if (kind == PULSE_AGENT_EXTERNAL) continue;
    uint32_t w0, w1;
    if (chance2) { w0 = chance2[2 * g]; w1 = chance2[2 * g + 1]; }
    else { PulseU4 b = pulse_tape(seed, env_base + (uint64_t)g, PULSE_STREAM_ACTION, counter); w0 = b.x; w1 = b.y; }
    const float* row = obs + g * W;
    int c0 = 0, c1 = 0;
    float pot = 0.0f;
    if (kind != PULSE_AGENT_RANDOM) { c0 = (int)row[5]; c1 = (int)row[6]; }
    if (kind == PULSE_AGENT_SMALL_BALL) pot = row[9];
    actions[g] = pulse_policy_action(kind, c0, c1, pot, w0, w1);
  }
}
}  // namespace

extern "C" int pulse_poker_policy(const float* obs, int obs_width, const int32_t* seat_idx, const uint8_t* seat_kinds,
                                  int n_seats, const uint32_t* chance2_or_null, uint64_t seed, uint64_t env_base,
                                  uint32_t counter, int64_t* actions_inout, long long n_games, cudaStream_t stream) {
  PULSE_REQUIRE(n_games >= 0, "n_games < 0");
  PULSE_REQUIRE(n_seats >= 1 && n_seats <= 16, "n_seats must be in 1..16");
  PULSE_REQUIRE(obs_width >= 13, "obs_width < 13");
  PULSE_REQUIRE(seat_kinds != nullptr, "seat_kinds is null");
  PulseSeatKinds k;
  for (int s = 0; s < 16; ++s) k.kind[s] = s < n_seats ? seat_kinds[s] : (uint8_t)PULSE_AGENT_EXTERNAL;
  for (int s = 0; s < n_seats; ++s) PULSE_REQUIRE(k.kind[s] <= PULSE_AGENT_SMALL_BALL, "unknown agent kind");
  if (n_games == 0) return PULSE_OK;
  PULSE_REQUIRE(obs && seat_idx && actions_inout, "null buffer");
  poker_policy_kernel<<<pulse_grid_for(n_games, kBlock), kBlock, 0, stream>>>(
      obs, obs_width, seat_idx, k, chance2_or_null, seed, env_base, counter, actions_inout, n_games);
  return pulse_check_launch("pulse_poker_policy");
}

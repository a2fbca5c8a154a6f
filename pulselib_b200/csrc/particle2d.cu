// Particle2D: batched 2-D point mass, fp32 semi-implicit Euler.
// Behaviour follows /root/reference/environments/Particle2D/Particle2D.py:15-30 (rows D1, D2 of
// SURVEY.md 8(a)): clamp action to [-1,1]; vel += a*dt; pos += vel*dt (updated velocity);
// dist = ||pos||; reward = -dist - 0.001*sum(a^2); steps += 1;
// terminated = (dist < 0.1) | (steps >= max_steps) -- recomputed each step, never sticky, and
// terminated envs keep integrating.
//
// One thread per env.  state is [N,4] f32 so each thread moves one float4 in and out
// (16 B/lane, fully coalesced); obs is the reference's `state.clone()`.  HBM-bound:
// 70 algorithmic bytes per env-step.
#include "common.cuh"

// Products are rounded to fp32 before the add (__fmul_rn/__fadd_rn are never contracted into
// FMA), matching torch's separate mul and add kernels.
__global__ void __launch_bounds__(256)
particle2d_step_kernel(float4* __restrict__ state, const float2* __restrict__ action,
                       int32_t* __restrict__ steps, float4* __restrict__ obs_out,
                       float* __restrict__ reward_out, uint8_t* __restrict__ terminated_out,
                       uint8_t* __restrict__ truncated_out, float dt, int max_steps, int n) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    float4 s = state[i];
    float2 a = action[i];
    a.x = fminf(fmaxf(a.x, -1.0f), 1.0f);
    a.y = fminf(fmaxf(a.y, -1.0f), 1.0f);
    s.z = __fadd_rn(s.z, __fmul_rn(a.x, dt));
    s.w = __fadd_rn(s.w, __fmul_rn(a.y, dt));
    s.x = __fadd_rn(s.x, __fmul_rn(s.z, dt));
    s.y = __fadd_rn(s.y, __fmul_rn(s.w, dt));
    float dist = __fsqrt_rn(__fadd_rn(__fmul_rn(s.x, s.x), __fmul_rn(s.y, s.y)));
    float a2 = __fadd_rn(__fmul_rn(a.x, a.x), __fmul_rn(a.y, a.y));
    float r = __fsub_rn(-dist, __fmul_rn(0.001f, a2));
    int st = steps[i] + 1;
    state[i] = s;
    obs_out[i] = s;
    steps[i] = st;
    reward_out[i] = r;
    terminated_out[i] = (uint8_t)((dist < 0.1f) | (st >= max_steps));
    if (truncated_out) truncated_out[i] = 0;
  }
}

// reset: pos = 5*N(0,1) from the chance tape (Box-Muller on two draws), vel = 0, steps = 0.
// The reference draws torch.randn (Particle2D.py:16); the tape replaces that generator and the
// parity harness injects the same normals into the reference through env.state.
__global__ void __launch_bounds__(256)
particle2d_reset_kernel(float4* __restrict__ state, int32_t* __restrict__ steps,
                        uint8_t* __restrict__ terminated, float4* __restrict__ obs_out,
                        uint64_t seed, uint64_t env_base, uint32_t episode, int n) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    PulseU4 u = pulse_tape(seed, env_base + (uint64_t)i, PULSE_STREAM_P2D_INIT, episode);
    // u1 in (0,1], u2 in [0,1)
    float u1 = __fmul_rn((float)((u.x >> 8) + 1u), 5.9604644775390625e-08f);
    float u2 = __fmul_rn((float)(u.y >> 8), 5.9604644775390625e-08f);
    float rad = __fsqrt_rn(__fmul_rn(-2.0f, logf(u1)));
    float ang = __fmul_rn(6.2831853071795864769f, u2);
    float4 s;
    s.x = __fmul_rn(__fmul_rn(rad, cosf(ang)), 5.0f);
    s.y = __fmul_rn(__fmul_rn(rad, sinf(ang)), 5.0f);
    s.z = 0.0f; s.w = 0.0f;
    state[i] = s;
    obs_out[i] = s;
    steps[i] = 0;
    terminated[i] = 0;
  }
}

extern "C" {

int pulse_particle2d_step(float* state, const float* action, int32_t* steps, float* obs_out,
                          float* reward_out, uint8_t* terminated_out, uint8_t* truncated_out_or_null,
                          float dt, int max_steps, int n_envs, cudaStream_t stream) {
  PULSE_REQUIRE(n_envs >= 0, "n_envs < 0");
  if (n_envs == 0) return PULSE_OK;
  PULSE_REQUIRE(state && action && steps && obs_out && reward_out && terminated_out, "null buffer");
  const int block = 256;
  particle2d_step_kernel<<<pulse_grid_for(n_envs, block), block, 0, stream>>>(
      (float4*)state, (const float2*)action, steps, (float4*)obs_out, reward_out, terminated_out,
      truncated_out_or_null, dt, max_steps, n_envs);
  return pulse_check_launch("pulse_particle2d_step");
}

int pulse_particle2d_reset(float* state, int32_t* steps, uint8_t* terminated, float* obs_out,
                           uint64_t seed, uint64_t env_base, uint32_t episode, int n_envs,
                           cudaStream_t stream) {
  PULSE_REQUIRE(n_envs >= 0, "n_envs < 0");
  if (n_envs == 0) return PULSE_OK;
  PULSE_REQUIRE(state && steps && terminated && obs_out, "null buffer");
  const int block = 256;
  particle2d_reset_kernel<<<pulse_grid_for(n_envs, block), block, 0, stream>>>(
      (float4*)state, steps, terminated, (float4*)obs_out, seed, env_base, episode, n_envs);
  return pulse_check_launch("pulse_particle2d_reset");
}
}

// TEST INFRASTRUCTURE ONLY.  Compiles the per-env bodies of the CUDA kernels (the __host__
// __device__ functions in pulselib_b200/csrc/*_body.cuh) for the CPU and loops them over the batch,
// so the CPU test suite can check the kernel logic against the oracle without a GPU.  Not part of
// libpulse_envs.so and never used by the product.
#include "poker_compat_body.cuh"
#include "poker_packed_body.cuh"
#include "blackjack_body.cuh"
#include "tfe_body.cuh"
#include "tetris_body.cuh"

void pulse_set_error(const char*, ...) {}
int pulse_check_launch(const char*) { return 0; }
int pulse_sm_count(void) { return 148; }

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
// scripted opponents: the per-table rule on observation rows, and fused into the packed step
int emu_poker_policy(const float* obs, int W, const int32_t* seat_idx, const uint8_t* kinds, int n_seats,
                     const uint32_t* chance2, uint64_t seed, uint64_t env_base, uint32_t counter, int64_t* actions,
                     long long n) {
  for (long long g = 0; g < n; ++g) {
    const int seat = seat_idx[g];
    const int kind = (seat >= 0 && seat < n_seats) ? kinds[seat] : PULSE_AGENT_EXTERNAL;
    if (kind == PULSE_AGENT_EXTERNAL) continue;
    uint32_t w0, w1;
    if (chance2) { w0 = chance2[2 * g]; w1 = chance2[2 * g + 1]; }
    else { PulseU4 b = pulse_tape(seed, env_base + (uint64_t)g, PULSE_STREAM_ACTION, counter); w0 = b.x; w1 = b.y; }
    const float* row = obs + g * W;
    actions[g] = pulse_policy_action(kind, (int)row[5], (int)row[6], row[9], w0, w1);
  }
  return 0;
}
int emu_poker_packed_step_policy(const PulsePokerPacked* st, const PulsePokerConfig* cfg, const int64_t* actions,
                                 const uint8_t* kinds, const uint32_t* chance2, uint32_t counter, int64_t* chosen,
                                 float* obs_out, float* reward_out, uint8_t* done_out, int auto_reset, uint64_t seed,
                                 uint64_t env_base) {
  PokerParams pr = mk(cfg);
  const int W = 13 + 3 * (pr.M - 1);
  PackedPolicy pol;
  for (int s = 0; s < 16; ++s) pol.kinds.kind[s] = s < cfg->n_players ? kinds[s] : (uint8_t)PULSE_AGENT_EXTERNAL;
  pol.chance2 = chance2; pol.chosen = chosen; pol.seat_before = nullptr; pol.counter = counter;
  for (long long g = 0; g < cfg->n_games; ++g)
    EMU_DISPATCH(pr.P, (packed_step_one<MP, true>(*st, pr, actions, obs_out + g * W, reward_out, done_out, auto_reset,
                                                  seed, env_base, cfg->starting_bbs, cfg->max_bbs, g,
                                                  (long long)cfg->n_games, &pol)));
  return 0;
}
// the compact step record (poker_wire.cuh) of every table of a packed batch, reward taken from `rewards`
int emu_poker_wire_rows(const PulsePokerPacked* st, const PulsePokerConfig* cfg, const float* rewards, uint32_t* rows) {
  PokerParams pr = mk(cfg);
  const long long n = cfg->n_games;
#define EMU_WIRE(MPV, MTV)                                                         \
  for (long long g = 0; g < n; ++g) {                                              \
    PokerTable<MPV> t; uint32_t fb;                                                \
    packed_load(t, fb, *st, g, n, pr.AP);                                          \
    uint32_t w[pkw_row_words(MTV)];                                                \
    pkw_pack_row<MPV, MTV>(t, pr, rewards[g], w);                                  \
    for (int i = 0; i < pkw_row_words(MTV); ++i) rows[g * pkw_row_words(MTV) + i] = w[i]; \
  }
  if (pr.P <= 2 && pr.M == 2) { EMU_WIRE(2, 2) }
  else if (pr.P <= 4 && pr.M == 4) { EMU_WIRE(4, 4) }
  else if (pr.P <= 6 && pr.M == 6) { EMU_WIRE(6, 6) }
  else if (pr.M == 10) { EMU_WIRE(10, 10) }
  else return -1;
#undef EMU_WIRE
  return 0;
}
float emu_enumerated_strength(int stage, int h0, int h1, uint32_t fb) { return packed_enumerated_strength(stage, h0, h1, fb); }
// showdown equity of every table, the boards of a table split over `nl` emulated lanes (1 = serial, 32 = as the kernel's warp)
int emu_showdown_equity(const PulsePokerPacked* st, const PulsePokerConfig* cfg, float* out, int samples, uint64_t seed,
                        uint64_t env_base, uint32_t counter, int nl) {
  PokerParams pr = mk(cfg);
  const long long n = cfg->n_games;
#define EMU_EQ(MPV)                                                                                          \
  for (long long g = 0; g < n; ++g) {                                                                        \
    PokerTable<MPV> t; uint32_t fb;                                                                          \
    packed_load(t, fb, *st, g, n, pr.AP);                                                                    \
    int64_t tot[MPV] = {0};                                                                                  \
    int boards = 0;                                                                                          \
    for (int lane = 0; lane < nl; ++lane) {                                                                  \
      int32_t num[MPV] = {0};                                                                                \
      boards = packed_showdown_equity_part<MPV>(t, pr.AP, samples, seed, env_base + (uint64_t)g, counter, lane, nl, num); \
      for (int s = 0; s < MPV; ++s) tot[s] += num[s];                                                        \
    }                                                                                                        \
    for (int s = 0; s < pr.AP; ++s) out[g * pr.AP + s] = pkq_equity(tot[s], boards);                         \
  }
  if (pr.P <= 2) { EMU_EQ(2) } else if (pr.P <= 4) { EMU_EQ(4) } else if (pr.P <= 6) { EMU_EQ(6) } else { EMU_EQ(10) }
#undef EMU_EQ
  return 0;
}
int emu_poker_pack(const PulsePokerState* compat, const PulsePokerPacked* st, const PulsePokerConfig* cfg) {
  PokerParams pr = mk(cfg);
  int bad = 0;
  for (long long g = 0; g < cfg->n_games; ++g) {
    bool ok = true;
    EMU_DISPATCH(pr.P, (ok = packed_pack_one<MP>(*st, *compat, pr, g, (long long)cfg->n_games)));
    bad += ok ? 0 : 1;
  }
  return bad;
}
int emu_poker_unpack(const PulsePokerPacked* st, const PulsePokerState* compat, const PulsePokerConfig* cfg) {
  PokerParams pr = mk(cfg);
  for (long long g = 0; g < cfg->n_games; ++g)
    EMU_DISPATCH(pr.P, (packed_unpack_one<MP>(*st, *compat, pr, g, (long long)cfg->n_games)));
  return 0;
}
int emu_poker_unpack_full(const PulsePokerPacked* st, const PulsePokerState* compat, const PulsePokerConfig* cfg,
                          const int32_t* prefixed, uint64_t seed, uint64_t env_base, int fresh) {
  PokerParams pr = mk(cfg);
  for (long long g = 0; g < cfg->n_games; ++g)
    EMU_DISPATCH(pr.P, (packed_unpack_one<MP>(*st, *compat, pr, g, (long long)cfg->n_games, true, prefixed, seed,
                                              env_base, fresh != 0)));
  return 0;
}

// the computed evaluator (csrc/poker_eval.cuh) against a HandRanks table walk over every stride-th n-card hand in
// lexicographic order; returns the number of disagreements, *checked = hands compared
long long emu_eval_vs_table(const int32_t* hr, int n_cards, int stride, long long* checked) {
  int c[7];
  long long bad = 0, seen = 0, idx = 0;
  for (int i = 0; i < n_cards; ++i) c[i] = i + 1;
  for (;;) {
    if (idx % stride == 0) {
      uint64_t m = 0;
      int p = 53;
      for (int i = 0; i < n_cards; ++i) { m = pe_add_card64(c[i], m); p = hr[p + c[i]]; }
      if (n_cards < 7) p = hr[p];
      bad += pe_eval_mask64(m) != p;
      ++seen;
    }
    ++idx;
    int k = n_cards - 1;
    while (k >= 0 && c[k] == 52 - (n_cards - 1 - k)) --k;
    if (k < 0) break;
    ++c[k];
    for (int j = k + 1; j < n_cards; ++j) c[j] = c[j - 1] + 1;
  }
  if (checked) *checked = seen;
  return bad;
}

// ---- Blackjack ----
int emu_blackjack_reset(uint32_t* st, uint8_t* deck, const uint8_t* injected, uint64_t seed, uint64_t env_base,
                        uint32_t episode, int32_t* obs, int32_t* rewards, uint8_t* terminated, int n) {
  for (long long g = 0; g < n; ++g) bj_reset_one(st, deck, injected, seed, env_base, episode, obs, rewards, terminated, g, n);
  return 0;
}
int emu_blackjack_step(uint32_t* st, const uint8_t* deck, const int64_t* actions, int32_t* obs, int32_t* rewards,
                       uint8_t* terminated, int n) {
  for (long long g = 0; g < n; ++g) bj_step_one(st, deck, actions, obs, rewards, terminated, g, n);
  return 0;
}
int emu_blackjack_rollout(uint32_t* st, uint8_t* deck, const int64_t* actions, int K, int32_t* obs, int32_t* rewards,
                          uint8_t* terminated, int n) {
  for (long long g = 0; g < n; ++g) {
    BjState s = bj_unpack_word(st[g]);
    for (int k = 0; k < K; ++k) {
      const long long o = (long long)k * n + g;
      rewards[o] = bj_step_state(s, actions[o], deck, g, n);
      obs[3 * o] = s.psum; obs[3 * o + 1] = s.ace; obs[3 * o + 2] = s.upcard; terminated[o] = s.term;
    }
    st[g] = bj_pack(s);
  }
  return 0;
}
int emu_blackjack_unpack(const uint32_t* st, const uint8_t* deck, const BjCompat* C, int n) {
  for (long long g = 0; g < n; ++g) bj_unpack_one(st, deck, *C, g, n);
  return 0;
}

}  // extern "C"

// ---- 2048 / TFE ----
template <int SIDE>
static void emu_tfe_step_t(uint64_t* boards, const int64_t* actions, const uint32_t* chance, const uint32_t* reset_chance,
                           uint64_t seed, uint64_t env_base, uint32_t counter, int auto_reset, int32_t* obs,
                           int32_t* reward, uint8_t* done, int64_t* score, int n) {
  for (long long g = 0; g < n; ++g) {
    const uint64_t b = boards[g];
    const bool was_done = auto_reset && tfe_game_over<SIDE>(b);
    PulseU4 sp, rc;
    if (chance) { sp.x = chance[2 * g]; sp.y = chance[2 * g + 1]; sp.z = sp.w = 0; }
    else sp = pulse_tape(seed, env_base + (uint64_t)g, PULSE_STREAM_TFE_SPAWN, counter);
    if (reset_chance) { rc.x = reset_chance[4 * g]; rc.y = reset_chance[4 * g + 1]; rc.z = reset_chance[4 * g + 2]; rc.w = reset_chance[4 * g + 3]; }
    else rc = pulse_tape(seed, env_base + (uint64_t)g, PULSE_STREAM_TFE_RESET, counter);
    int32_t r; bool d;
    boards[g] = tfe_step_one<SIDE>(b, (int)actions[g], was_done, auto_reset, sp, rc, score[g], r, d, TfeComputedRows());
    reward[g] = r; done[g] = d;
    if (obs) tfe_unpack_board<SIDE>(boards[g], obs + g * SIDE * SIDE);
  }
}
template <int SIDE>
static void emu_tfe_reset_t(uint64_t* boards, int64_t* score, uint8_t* done, const uint32_t* chance, uint64_t seed,
                            uint64_t env_base, uint32_t counter, int32_t* obs, int n) {
  for (long long g = 0; g < n; ++g) {
    PulseU4 c;
    if (chance) { c.x = chance[4 * g]; c.y = chance[4 * g + 1]; c.z = chance[4 * g + 2]; c.w = chance[4 * g + 3]; }
    else c = pulse_tape(seed, env_base + (uint64_t)g, PULSE_STREAM_TFE_RESET, counter);
    boards[g] = tfe_reset_board<SIDE>(c); score[g] = 0; done[g] = 0;
    if (obs) tfe_unpack_board<SIDE>(boards[g], obs + g * SIDE * SIDE);
  }
}
extern "C" {
int emu_tfe_step(uint64_t* boards, const int64_t* actions, const uint32_t* chance, const uint32_t* reset_chance,
                 uint64_t seed, uint64_t env_base, uint32_t counter, int auto_reset, int32_t* obs, int32_t* reward,
                 uint8_t* done, int64_t* score, int n, int side) {
  if (side == 4) emu_tfe_step_t<4>(boards, actions, chance, reset_chance, seed, env_base, counter, auto_reset, obs, reward, done, score, n);
  else if (side == 3) emu_tfe_step_t<3>(boards, actions, chance, reset_chance, seed, env_base, counter, auto_reset, obs, reward, done, score, n);
  else emu_tfe_step_t<2>(boards, actions, chance, reset_chance, seed, env_base, counter, auto_reset, obs, reward, done, score, n);
  return 0;
}
int emu_tfe_reset(uint64_t* boards, int64_t* score, uint8_t* done, const uint32_t* chance, uint64_t seed, uint64_t env_base,
                  uint32_t counter, int32_t* obs, int n, int side) {
  if (side == 4) emu_tfe_reset_t<4>(boards, score, done, chance, seed, env_base, counter, obs, n);
  else if (side == 3) emu_tfe_reset_t<3>(boards, score, done, chance, seed, env_base, counter, obs, n);
  else emu_tfe_reset_t<2>(boards, score, done, chance, seed, env_base, counter, obs, n);
  return 0;
}
// the row-move table as the device builds it, exercised through TfeTableRows on every 16-bit row
int emu_tfe_table_check(void) {
  static uint16_t res[65536];
  static uint8_t code[65536];
  for (uint32_t r = 0; r < 65536; ++r) { uint32_t sc = 0; res[tfe_row_slot(r)] = (uint16_t)tfe_squash_row(r, sc); code[tfe_row_slot(r)] = (uint8_t)tfe_score_code(r); }
  static uint32_t sc_of[256];
  for (uint32_t c = 0; c < 256; ++c) sc_of[c] = tfe_code_score(c);
  TfeTableRows t{res, code, sc_of};
  int bad = 0;
  for (uint32_t r = 0; r < 65536; ++r) {
    uint32_t s1 = 0, s2 = 0;
    const uint32_t a = tfe_squash_row(r, s1), b = t(r, s2);
    bad += (a != b) || (s1 != s2);
  }
  return bad;
}
// tfe_board_slots (the four rows of a board hashed at once) against tfe_row_slot row by row, and tfe_row_slot
// as a bijection on 16 bits; returns the number of disagreements
int emu_tfe_slot_check(const uint64_t* boards, int n) {
  int bad = 0;
  for (int i = 0; i < n; ++i) {
    const uint64_t s = tfe_board_slots(boards[i]);
    for (int r = 0; r < 4; ++r)
      bad += (uint32_t)((s >> (16 * r)) & 0xFFFFu) != tfe_row_slot((uint32_t)((boards[i] >> (16 * r)) & 0xFFFFu));
  }
  static uint8_t seen[65536];
  for (uint32_t r = 0; r < 65536; ++r) seen[r] = 0;
  for (uint32_t r = 0; r < 65536; ++r) { const uint32_t t = tfe_row_slot(r); bad += t > 0xFFFFu || seen[t]; seen[t & 0xFFFFu] = 1; }
  return bad;
}
int emu_tfe_pack(const int32_t* values, uint64_t* boards, int n, int side) {
  int bad = 0;
  for (long long g = 0; g < n; ++g) {
    bool ok = side == 4 ? tfe_pack_board<4>(values + g * 16, boards[g]) : side == 3 ? tfe_pack_board<3>(values + g * 9, boards[g])
                                                                                    : tfe_pack_board<2>(values + g * 4, boards[g]);
    bad += !ok;
  }
  return bad;
}

// ---- Tetris ----
int emu_tetris_step(uint16_t* rows, uint8_t* piece, int32_t* lines, uint8_t* done, const int64_t* actions,
                    const uint8_t* next_piece, const uint8_t* reset_piece, uint64_t seed, uint64_t env_base,
                    uint32_t counter, int auto_reset, int32_t* reward, int n) {
  static const uint32_t shapes[TET_SHAPES] = TET_SHAPE_TABLE;
  for (long long g = 0; g < n; ++g) {
    const PulseU4 blk = pulse_tape(seed, env_base + (uint64_t)g, PULSE_STREAM_TETRIS_PIECE, counter);
    const int nx = next_piece ? next_piece[g] : (int)pulse_mulhi32(blk.x, 7u);
    const int rp = reset_piece ? reset_piece[g] : (int)pulse_mulhi32(blk.y, 7u);
    int pc = piece[g]; int32_t ln = lines[g]; bool dn = done[g]; int32_t r;
    tet_step_one(rows + g * TET_H, shapes, pc, ln, dn, actions[g], nx, rp, auto_reset, r);
    piece[g] = (uint8_t)pc; lines[g] = ln; done[g] = dn; reward[g] = r;
  }
  return 0;
}
}

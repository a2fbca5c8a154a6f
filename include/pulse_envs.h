/* libpulse_envs -- C ABI of the B200-native batched env-step engine.
 *
 * The reference (cerredz/Pulselib) has no FFI: its env step is Python (torch eager / numba).  The
 * boundary this library sits behind is therefore the Python gym.Env class contract; each entry
 * point below names the reference method it replaces.  pulselib_b200/envs/ binds these with ctypes
 * (see INTEGRATION.md for the stub a Pulselib maintainer would add).
 *
 * Conventions for every call:
 *   - returns 0 on success, a negative PULSE_E* code for bad arguments, or a positive cudaError_t;
 *     pulse_last_error() then holds a thread-local message;
 *   - never allocates, frees or retains caller buffers; never synchronises the stream;
 *   - unless a parameter says "host", pointers are device pointers the caller keeps alive until the
 *     stream has run the call; `stream` is a cudaStream_t (e.g. torch.cuda.current_stream().cuda_stream);
 *   - the caller's current CUDA device must be the one owning the buffers;
 *   - thread-compatible: concurrent calls on different envs/streams are fine.
 */
#ifndef PULSE_ENVS_H
#define PULSE_ENVS_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct CUstream_st* pulse_stream_t; /* == cudaStream_t */

#define PULSE_OK 0
#define PULSE_EINVAL (-1)
#define PULSE_ENOTREADY (-2)
#define PULSE_EUNSUPPORTED (-3)

int pulse_version(void);
const char* pulse_last_error(void);

/* ---- chance tape ------------------------------------------------------------------------------
 * Philox-4x32-10, counter {env_id lo, env_id hi, draw counter, stream tag}, key = seed.  Host
 * helper so callers can materialise the draws the kernels consume. */
int pulse_tape_block(uint64_t seed, uint64_t env_id, uint32_t stream_tag, uint32_t counter, uint32_t* host_out4);

/* ---- HandRanks table (replaces the HandRanks.dat download, PokerGPU.py:47-58) ------------------ */
size_t pulse_handranks_entries(void);                                   /* 32,487,834 */
int pulse_handranks_generate(int32_t* host_out, size_t n_entries);      /* host, ~6 s */
int pulse_handranks_walk(const int32_t* host_table, const int32_t* host_cards, int n_cards);
int pulse_eval_cards(const int32_t* host_cards, int n_cards);           /* computed evaluator, host */
/* computed evaluator on the device: cards int32 [n, n_cards] (5..7), out int32 [n] */
int pulse_poker_eval_batch(const int32_t* cards, int n_cards, int32_t* out, int n, pulse_stream_t stream);

/* ---- Particle2D (environments/Particle2D/Particle2D.py:15-30) --------------------------------- */
int pulse_particle2d_reset(float* state /*[N,4]*/, int32_t* steps, uint8_t* terminated, float* obs_out /*[N,4]*/,
                           uint64_t seed, uint64_t env_base, uint32_t episode, int n_envs, pulse_stream_t stream);
int pulse_particle2d_step(float* state /*[N,4]*/, const float* action /*[N,2]*/, int32_t* steps,
                          float* obs_out /*[N,4]*/, float* reward_out, uint8_t* terminated_out,
                          uint8_t* truncated_out_or_null, float dt, int max_steps, int n_envs,
                          pulse_stream_t stream);

/* ---- 2048 / TFE (environments/2048/TFE.py:17-189), batched --------------------------------------
 * boards uint64 [N]: tile exponents, 4 bits per cell, cell (r, c) at bits 16 r + 4 c (side 2..4,
 * square).  score int64 [N] = TFE.total_score.  obs (optional) int32 [N, side, side] tile values, the
 * reference's board representation (:121).  Chance: add_tile_numba's two RNG draws (:31-33) are two
 * 32-bit words per spawn, idx = mulhi32(u0, n_empty), four = u1 > floor(0.9 * 2^32); injected as
 * chance2 uint32 [N,2] (reset: chance4 uint32 [N,4] = two spawns) or, when NULL, taken from the tape
 * at (seed, env_base + i, counter) -- pass a different counter for every reset / step.
 * reset = TFE.reset (:143-149).  step = TFE.step (:152-189): action 0 left, 1 up, 2 right, 3 down
 * (uint8 or int64, action_bytes = 1 or 8); reward = floor(log2(step score)) or 0 (:185-187); done =
 * is_game_over_numba (:48-67).  auto_reset != 0 (engine extension; the reference has no batched
 * form): a board whose done flag is set when the step begins (done_inout, as left by the previous
 * reset / step / pack) is reset instead (reward 0, done 0).  obs must be 16-byte aligned.
 * counter_dev_or_null: uint32 [2] on the device, zero-initialised by the caller; when given, the tape
 * counter of this launch is counter + counter_dev[0] and the launch increments counter_dev[0] when it
 * finishes, so one captured launch can be replayed from a CUDA graph with advancing chance. */
/* Builds this device's row-move table (192 KB, staged in shared memory by every step launch) if it does
 * not exist yet: one cudaMalloc and one stream synchronisation, once per device.  pulse_tfe_step does the
 * same on first use; call this first when the first step would fall inside a stream capture. */
int pulse_tfe_init(pulse_stream_t stream);
int pulse_tfe_reset(uint64_t* boards, int64_t* score, uint8_t* done, const uint32_t* chance4_or_null, uint64_t seed,
                    uint64_t env_base, uint32_t counter, int32_t* obs_out_or_null, int n_envs, int side,
                    pulse_stream_t stream);
int pulse_tfe_step(uint64_t* boards, const void* actions, int action_bytes, const uint32_t* chance2_or_null,
                   const uint32_t* reset_chance4_or_null, uint64_t seed, uint64_t env_base, uint32_t counter,
                   uint32_t* counter_dev_or_null, int auto_reset, int32_t* obs_out_or_null, int32_t* reward_out,
                   uint8_t* done_inout, int64_t* score_inout, int n_envs, int side, pulse_stream_t stream);
/* K steps per launch (SURVEY.md 8(b) "K steps per launch"), for open-loop action tapes: actions uint8 / int64 [K, N];
 * outputs step-major: obs_out int32 [K, N, side, side] (or NULL), reward_out int32 [K, N], done_out uint8 [K, N] --
 * what K calls of pulse_tfe_step with tape counters counter .. counter + K - 1 would have written one after another
 * (chance from the tape only).  The row table is staged once per K steps and a board stays in registers between its
 * steps.  done_inout uint8 [N]: the per-board done flag as pulse_tfe_step keeps it. */
int pulse_tfe_rollout(uint64_t* boards, const void* actions, int action_bytes, int n_steps, uint64_t seed,
                      uint64_t env_base, uint32_t counter, int auto_reset, int32_t* obs_out_or_null, int32_t* reward_out,
                      uint8_t* done_out, uint8_t* done_inout, int64_t* score_inout, int n_envs, int side,
                      pulse_stream_t stream);
int pulse_tfe_unpack(const uint64_t* boards, int32_t* values_out, int n_envs, int side, pulse_stream_t stream);
/* Domain audit of the packed boards (tile exponents up to 15 = 32768; two 32768 tiles meeting stay unmerged where the
 * reference would make 65536).  counts2 int [2] on the device, accumulated: [0] += boards holding a 32768 tile,
 * [1] += boards holding two or more (one move away from leaving the packed domain).  Random play never gets near
 * (a 32768 tile takes > 10^4 moves); long-running learned policies should poll it. */
int pulse_tfe_domain_check(const uint64_t* boards, int* counts2, int n_envs, pulse_stream_t stream);
/* int32 tile values -> packed boards (+ their game-over flags); *bad_count_or_null is incremented for
 * every board holding a value that is not 0 or a power of two in 2..32768 */
int pulse_tfe_pack(const int32_t* values, uint64_t* boards_out, uint8_t* done_out_or_null, int* bad_count_or_null,
                   int n_envs, int side, pulse_stream_t stream);

/* ---- Tetris 10x20, batched placement + line clear ------------------------------------------------
 * The reference ships only empty shells (environments/Tetris/Tetris.py:1-13, TetrisNumba.py:1-16), so
 * these entry points replace no reference code; the rules are DESIGN.md section 6 / oracle/oracle_tetris.c.
 * rows uint16 [N,20] row-major, 16-byte aligned (bit x of rows[i][y] = cell (x, y), row 0 bottom; this
 * tensor is also the observation); piece uint8 [N] current piece 0..6 (I O T S Z J L); lines int32 [N];
 * done uint8/bool [N].  action in 0..39 = rotation * 10 + column (uint8 or int64).  Chance: the piece
 * that follows a placement (next_piece uint8 [N]) and the first piece after an auto-reset (reset_piece
 * uint8 [N]); NULL takes mulhi32(word 0 / word 1, 7) of the tape block at (seed, env_base + i, counter).
 * reward int32 [N] = rows cleared by this step.  auto_reset != 0: an env that is done when the step
 * begins is cleared instead of stepped (reward 0, done 0).  counter_dev_or_null: as for pulse_tfe_step. */
int pulse_tetris_reset(uint16_t* rows, uint8_t* piece, int32_t* lines, uint8_t* done, const uint8_t* first_piece_or_null,
                       uint64_t seed, uint64_t env_base, uint32_t counter, int n_envs, pulse_stream_t stream);
int pulse_tetris_step(uint16_t* rows, uint8_t* piece, int32_t* lines, uint8_t* done, const void* actions,
                      int action_bytes, const uint8_t* next_piece_or_null, const uint8_t* reset_piece_or_null,
                      uint64_t seed, uint64_t env_base, uint32_t counter, uint32_t* counter_dev_or_null,
                      int auto_reset, int32_t* reward_out, int n_envs, pulse_stream_t stream);

/* ---- Blackjack (environments/blackjack/blackjack.py:23-186) ------------------------------------
 * Engine state: st uint32 [N] (packed sums, counters, flags; layout in csrc/blackjack_body.cuh) and
 * deck uint8 [52][N] (position-major).  obs int32 [N,3] = (player_sum, usable_ace, dealer_upcard)
 * (:103-108); rewards int32 [N] in {-1,0,1} (:164-177); terminated uint8/bool [N].
 * reset = BlackJack.reset + deal_starting_cards (:23-101); injected_decks_or_null: uint8 [N,52]
 * row-major permutations of 0..51, NULL draws them from the chance tape (seed, env_base+i, episode).
 * step = BlackJack.step (:179-186): action 0 hit, 1 stand, anything else / terminated env: no-op. */
int pulse_blackjack_reset(uint32_t* st, uint8_t* deck, const uint8_t* injected_decks_or_null, uint64_t seed,
                          uint64_t env_base, uint32_t episode, int32_t* obs_out, int32_t* rewards_out,
                          uint8_t* terminated_out, int n_envs, pulse_stream_t stream);
int pulse_blackjack_step(uint32_t* st, const uint8_t* deck, const int64_t* actions, int32_t* obs_out,
                         int32_t* rewards_out, uint8_t* terminated_out, int n_envs, pulse_stream_t stream);
/* K steps per launch: at the reference's own batch size (4096 envs) one step is launch-bound, so the
 * loop over steps runs inside the kernel with the env's state word in registers.  actions: uint8 or int64 [K, N]
 * (action_bytes 1 / 8); outputs step-major: obs_out int32 [K, N, 3], rewards_out int32 [K, N], terminated_out uint8
 * [K, N] -- exactly what K calls of pulse_blackjack_step would have written.  reset_first != 0: deal a fresh hand from
 * the tape (seed, env_base + i, episode) before the first step (= pulse_blackjack_reset; its observation goes to
 * reset_obs_out_or_null int32 [N, 3]).  n_hands > 1: that many hands of n_steps steps in the one launch, each after the
 * first (and the first too when reset_first) starting with a fresh deal from consecutive tape episodes; actions and
 * outputs are then [n_hands * n_steps, N, ...] and reset_obs_out [n_hands, N, 3]. */
int pulse_blackjack_rollout(uint32_t* st, uint8_t* deck, const void* actions, int action_bytes, int n_steps, int n_hands,
                            int reset_first, uint64_t seed, uint64_t env_base, uint32_t episode,
                            int32_t* reset_obs_out_or_null, int32_t* obs_out, int32_t* rewards_out,
                            uint8_t* terminated_out, int n_envs, pulse_stream_t stream);
/* engine state -> the reference's tensor attributes (blackjack.py:24-44), for inspection / parity */
typedef struct PulseBlackjackCompat {
  int64_t* decks;             /* [N,52] */
  int32_t* deck_positions;    /* [N] */
  int32_t* players_cards;     /* [N,20] */
  int32_t* players_card_idx;  /* [N] */
  int32_t* player_card_sums;  /* [N] */
  uint8_t* has_ace;           /* [N] */
  int32_t* dealer_cards;      /* [N,20] */
  int32_t* dealer_card_idx;   /* [N] */
  int32_t* dealer_upcard;     /* [N] */
  int32_t* dealer_card_sums;  /* [N] */
  uint8_t* dealer_has_ace;    /* [N] */
} PulseBlackjackCompat;
int pulse_blackjack_unpack(const uint32_t* st, const uint8_t* deck, const PulseBlackjackCompat* out, int n_envs,
                           pulse_stream_t stream);

/* ---- PokerGPU, reference-shaped layout (environments/Poker/PokerGPU.py) ----------------------- */
/* One device pointer per tensor attribute of the reference env, same shapes and dtypes (int32,
 * row-major; is_done / equity_dirty are torch.bool = 1 byte). */
typedef struct PulsePokerState {
  int32_t* decks;             /* [N,52] 1..52                        PokerGPU.py:86-92  */
  int32_t* deck_positions;    /* [N]                                 :93  */
  int32_t* board;             /* [N,5], -1 = not dealt               :95  */
  int32_t* pots;              /* [N]                                 :96  */
  int32_t* stages;            /* [N] 0..5                            :97  */
  int32_t* stacks;            /* [N,P]                               :101-110 */
  int32_t* hands;             /* [N,P,2]                             :112-114 */
  int32_t* current_round_bet; /* [N,P]                               :116 */
  int32_t* total_invested;    /* [N,P]                               :117 */
  int32_t* status;            /* [N,P] 0 ACTIVE 1 FOLDED 2 ALLIN 3 SITOUT  :118-119 */
  int32_t* button;            /* [N]                                 :121 */
  int32_t* sb;                /* [N]                                 :123-128 */
  int32_t* bb;                /* [N] */
  int32_t* idx;               /* [N] seat to act                     :130-133 */
  int32_t* highest;           /* [N]                                 :134 */
  int32_t* agg;               /* [N]                                 :135 */
  int32_t* acted;             /* [N]                                 :136 */
  int32_t* last_raise_size;   /* [N]                                 :81  */
  int32_t* prev_invested;     /* [N]                                 :151 */
  uint8_t* is_done;           /* [N]                                 :137 */
  uint8_t* equity_dirty;      /* [N]                                 :145 */
  float* equities;            /* [N,AP]                              :144 */
  const int32_t* hand_ranks;  /* HandRanks table on the device       :47-58 */
  int64_t n_hand_ranks;
} PulsePokerState;

typedef struct PulsePokerConfig {
  int32_t n_games, n_players, active_players, max_players;
  int32_t starting_bbs, max_bbs;
  int32_t K, alpha;
  float w1, w2;
} PulsePokerConfig;

/* reset (PokerGPU.py:73-157).  prefixed_decks_or_null: int32 [N,52] copied into st->decks; NULL
 * draws each deck from the chance tape (seed, env_base + table, hand_counter).  first_reset bit 0: stacks = starting_bbs
 * (else busted/oversized stacks are refilled and rolled by `rotation`); bit 1: button = 0 (else it
 * advances modulo the table size); bit 2 (packed layout only): restart the per-table hand index at hand_counter.
 * The first reset of an env passes 3.  Writes the first observation [N, 13+3(M-1)]. */
int pulse_poker_reset(const PulsePokerState* st, const PulsePokerConfig* cfg, const int32_t* prefixed_decks_or_null,
                      uint64_t seed, uint64_t env_base, uint32_t hand_counter, int rotation, int first_reset,
                      float* obs_out, pulse_stream_t stream);

/* Phases of a step, in the reference's order (PokerGPU.py:527-633). */
#define PULSE_POKER_EQUITIES 1        /* calculate_equities           :455-525 */
#define PULSE_POKER_EXECUTE 2         /* execute_actions              :230-303 */
#define PULSE_POKER_ADVANCE 4         /* next actor + street change   :547-616 */
#define PULSE_POKER_FOLD_WINNERS 8    /* resolve_fold_winners         :331-338 */
#define PULSE_POKER_SHOWDOWN 16       /* resolve_terminated_games     :380-453 */
#define PULSE_POKER_CLEAR_DONE 32     /*                              :625-628 */
#define PULSE_POKER_REWARD 64         /* poker_reward_gpu             :305-329 */
#define PULSE_POKER_OBS 128           /* get_obs                      :159-179 */
#define PULSE_POKER_STEP 255
#define PULSE_POKER_NEWLY_IS_DONE 256 /* resolve_* called directly: is_done is the mask */
#define PULSE_POKER_RAW_REWARD 512    /* reward for actor_idx[], not masked by legality */

/* step (PokerGPU.py:527-633) = one fused launch with phases = PULSE_POKER_STEP.  The reference's
 * internal methods are the same kernel with a single phase bit.  actions int64 [N]; done_out may
 * alias st->is_done (the reference rebinds is_done to a new tensor each step, :623).  Any of
 * obs_out / reward_out may be NULL when its phase is not requested. */
int pulse_poker_step(const PulsePokerState* st, const PulsePokerConfig* cfg, const int64_t* actions,
                     const int32_t* actor_idx_or_null, uint32_t phases, float* obs_out, float* reward_out,
                     uint8_t* done_out, pulse_stream_t stream);

/* ---- PokerGPU, packed layout (throughput path; layout in pulselib_b200/csrc/poker_packed_body.cuh) */
typedef struct PulsePokerPacked {
  uint32_t* tbl;             /* [N][4] meta | status planes | pot | highest */
  uint32_t* tbl2;            /* [N][2] last_raise_size | five board cards x 6 bits */
  int32_t* stack;            /* [P][N] */
  int32_t* crb;              /* [P][N] current_round_bet */
  int32_t* inv;              /* [P][N] total_invested */
  uint16_t* hole;            /* [P][N] card0 | card1 << 6 */
  uint64_t* ev;              /* [P][N] per-deal hand values: 7-card | 6-card << 16 | HR[5-card] << 32 */
  uint32_t* hand_ctr;        /* [N] hands dealt (needed for auto_reset), may be NULL otherwise */
  const int32_t* hand_ranks; /* HandRanks table (only its first 36,875 entries are read) */
  int64_t n_hand_ranks;
  /* optional live mirrors of two reference attributes, written by reset / step / redeal when non-NULL:
   * idx = the seat to act (info["seat_idx"], PokerGPU.py:181-186) and prev_invested (:539, :151) */
  int32_t* idx;              /* [N] or NULL */
  int32_t* prev_invested;    /* [N] or NULL */
} PulsePokerPacked;

/* Same contracts as pulse_poker_reset / pulse_poker_step.  auto_reset != 0 (engine extension): a
 * table already done when the step begins is dealt its next hand from the chance tape instead of
 * idling; that step returns reward 0, done 0 and the new hand's first observation.
 * The tape's hand index is monotonic per table: a reset deals hand max(hand_ctr[table] + 1, hand_counter) (first_reset
 * bit 2 set: exactly hand_counter, restarting the tape, e.g. after reseeding), a redeal hand_ctr[table] + 1 -- so
 * whole-batch resets mixed with per-table redeals never replay a deck. */
int pulse_poker_packed_reset(const PulsePokerPacked* st, const PulsePokerConfig* cfg,
                             const int32_t* prefixed_decks_or_null, uint64_t seed, uint64_t env_base,
                             uint32_t hand_counter, int rotation, int first_reset, float* obs_out_or_null,
                             pulse_stream_t stream);
int pulse_poker_packed_step(const PulsePokerPacked* st, const PulsePokerConfig* cfg, const int64_t* actions,
                            float* obs_out, float* reward_out, uint8_t* done_out, int auto_reset, uint64_t seed,
                            uint64_t env_base, pulse_stream_t stream);
/* ---- scripted opponents + build_actions ---------------------------------------
 * Replaces /root/reference/environments/Poker/utils.py:108-123 (build_actions) and the four scripted GPU
 * agents of /root/reference/environments/Poker/Player.py:79-176 with one rule per table.
 * seat_kinds: HOST array of n_seats bytes, one PULSE_AGENT_* per seat (the caller's rotated agent_types).
 * Chance: two 32-bit words per table (raise size = lo + mulhi32(w0, hi - lo), LoosePassive's uniform =
 * (w1 >> 8) * 2^-24), injected as chance2 uint32 [N,2] or taken from the tape at (seed, env_base + i,
 * stream 2, counter).  Tables whose acting seat is PULSE_AGENT_EXTERNAL keep the caller's action. */
enum {
  PULSE_AGENT_EXTERNAL = 0,         /* qlearning / host-side agent: action supplied by the caller */
  PULSE_AGENT_RANDOM = 1,           /* utils.py:119-120 */
  PULSE_AGENT_HEURISTIC_HANDS = 2,  /* Player.py:79-103 */
  PULSE_AGENT_TIGHT_AGGRESSIVE = 3, /* Player.py:105-126 */
  PULSE_AGENT_LOOSE_PASSIVE = 4,    /* Player.py:128-150 */
  PULSE_AGENT_SMALL_BALL = 5        /* Player.py:152-176 */
};
/* on the reference-shaped observation: obs float32 [N, obs_width], seat_idx int32 [N] (info["seat_idx"]),
 * actions_inout int64 [N] */
int pulse_poker_policy(const float* obs, int obs_width, const int32_t* seat_idx, const uint8_t* seat_kinds,
                       int n_seats, const uint32_t* chance2_or_null, uint64_t seed, uint64_t env_base,
                       uint32_t counter, int64_t* actions_inout, long long n_games, pulse_stream_t stream);
/* fused into the packed step: the acting seat's rule is evaluated on the registers the step already holds
 * (hole cards, pot), so a scripted seat costs no extra HBM traffic and no extra launch.  actions_or_null is
 * read only for PULSE_AGENT_EXTERNAL seats (may be NULL when every seat is scripted); chosen_out_or_null
 * int64 [N] receives the action that was played (what build_actions leaves in `actions`);
 * seat_idx_out_or_null int32 [N] the seat that acted (info["seat_idx"] as it was before the step). */
int pulse_poker_packed_step_policy(const PulsePokerPacked* st, const PulsePokerConfig* cfg,
                                   const int64_t* actions_or_null, const uint8_t* seat_kinds,
                                   const uint32_t* chance2_or_null, uint32_t counter, int64_t* chosen_out_or_null,
                                   int32_t* seat_idx_out_or_null, float* obs_out, float* reward_out, uint8_t* done_out, int auto_reset,
                                   uint64_t seed, uint64_t env_base, pulse_stream_t stream);

/* K steps per launch with every seat played on the device (SURVEY.md 8(b) "pulse_poker_rollout"): for the small batches
 * the reference benchmarks (N = 256, /root/reference/benchmarking/Poker) a step is launch-bound, so the loop over steps
 * runs inside one kernel.  seat_kinds: HOST array, one PULSE_AGENT_RANDOM .. PULSE_AGENT_SMALL_BALL per seat (no
 * EXTERNAL seats); step k draws its chance pair from the tape at counter + k, exactly as n_steps calls of
 * pulse_poker_packed_step_policy would.  Outputs step-major: reward_out float32 [K, N], done_out uint8 [K, N],
 * actions_out_or_null int64 [K, N] (the actions played); obs_final_or_null float32 [N, 13 + 3 (M - 1)] = the
 * observation after the last step. */
int pulse_poker_packed_rollout(const PulsePokerPacked* st, const PulsePokerConfig* cfg, const uint8_t* seat_kinds,
                               uint32_t counter, int n_steps, int auto_reset, uint64_t seed, uint64_t env_base,
                               float* obs_final_or_null, float* reward_out, uint8_t* done_out, int64_t* actions_out_or_null,
                               pulse_stream_t stream);
/* K steps per launch on SUPPLIED actions (open-loop tapes: replays, random-policy benchmarks): actions uint8 [K, N]
 * (255 = a negative id, which matches no branch).  Exactly what K calls of pulse_poker_packed_step without auto_reset
 * would return one after another -- reward_out float32 [K, N], done_out uint8 [K, N], obs_final_or_null = the
 * observation after the last step -- but the table is loaded once, stepped K times in registers and stored once. */
int pulse_poker_packed_rollout_actions(const PulsePokerPacked* st, const PulsePokerConfig* cfg, const uint8_t* actions,
                                       int n_steps, float* obs_final_or_null, float* reward_out, uint8_t* done_out,
                                       pulse_stream_t stream);

/* Enumerated hand strength (no reference behaviour -- the reference's notes call its table-walk
 * proxy a known weakness, rl/depth_notes.txt:302-333).  strength_out float32 [N, active_players]: the expectation over
 * the board cards still to come of the river strength clamp((v7 - 4109) / 32765, 0, 1) (PokerGPU.py:477-481) from each
 * seat's own point of view; river = the reference's value, turn = 46 cards, flop = 1081 pairs, preflop / finished
 * hands = 0.5.  An optional input for a caller's reward; the step's own reward keeps the reference's proxy. */
int pulse_poker_packed_enumerated_strength(const PulsePokerPacked* st, const PulsePokerConfig* cfg, float* strength_out,
                                           pulse_stream_t stream);
/* Showdown equity (second half; no reference behaviour -- rl/depth_notes.txt:316-321 asks for equity
 * "using ALL players card info" and calls it out of reach).  equity_out float32 [N, active_players]: the share of the
 * pot each contender (status ACTIVE or ALLIN) wins at showdown, averaged over every completion of the board from the
 * cards in nobody's hand and not on the visible board (folded seats' cards are dead); ties split; folded / sitting-out
 * seats get 0; a sole contender gets 1.  River and settled hands: the one board; turn and flop: exact enumeration;
 * preflop: preflop_samples (1..65536) boards drawn from the chance tape at (seed, env_base + table, counter, stream 8).
 * Shares are integer counts, so the result is independent of launch shape and bit-identical to oracle/oracle_core.c::
 * oracle_showdown_equity.  Rows of a table with at least one contender sum to 1 (within float rounding).  Reads the
 * packed state only. */
int pulse_poker_packed_showdown_equity(const PulsePokerPacked* st, const PulsePokerConfig* cfg, float* equity_out,
                                       int preflop_samples, uint64_t seed, uint64_t env_base, uint32_t counter,
                                       pulse_stream_t stream);
/* reference-shaped tensors -> packed (SURVEY.md 8(b) pulse_poker_pack): puts state that was edited through the
 * reference's attributes back onto the throughput layout.  Needs every state tensor of `compat` (equities,
 * equity_dirty, prev_invested, sb, bb optional).  The visible board comes from `board`, the cards still to come from
 * `decks`, the per-deal hand values are recomputed, hand_ctr is left alone.  *bad_count_or_null (device int) is
 * incremented for every table the packed layout cannot express exactly -- deck position other than the street's
 * (2 AP + 0 / 4 / 6 / 8), a board card beyond the street's visible ones, a card outside 1..52, seat / status / street
 * numbers out of range -- such a table is stored with the offending field forced into range and must not be trusted. */
int pulse_poker_pack(const PulsePokerState* compat, const PulsePokerPacked* st, const PulsePokerConfig* cfg,
                     int* bad_count_or_null, pulse_stream_t stream);
/* packed -> reference-shaped tensors.  Writes every field of `compat` whose pointer is non-NULL except decks
 * (a pure function of the deal, see pulse_poker_unpack_full) and prev_invested.  sb / bb follow from button and
 * active_players (PokerGPU.py:123-128); equities / equity_dirty are the reference's cache as it stands after the
 * last step: the values of the last street on which a step began (the packed table word keeps that street in two
 * bits), dirty where the street has changed since (:604-616). */
int pulse_poker_unpack(const PulsePokerPacked* st, const PulsePokerState* compat, const PulsePokerConfig* cfg,
                       pulse_stream_t stream);
/* The full mirror behind "Pulse-Poker-GPU-v1" (mirror on access): pulse_poker_unpack plus decks [N,52] -- copied
 * from prefixed_decks_or_null (the tensor the last reset was given) or re-drawn from the chance tape at
 * (seed, env_base + table, hand_ctr[table]) exactly as pulse_poker_reset draws them -- and prev_invested (from
 * st->prev_invested when the engine keeps it, else 0).  fresh != 0: no step has run since the last reset, every
 * table's equity cache is dirty (:145). */
int pulse_poker_unpack_full(const PulsePokerPacked* st, const PulsePokerState* compat, const PulsePokerConfig* cfg,
                            const int32_t* prefixed_decks_or_null, uint64_t seed, uint64_t env_base, int fresh,
                            pulse_stream_t stream);

/* ---- PokerGPU with HOST buffers (the end-to-end path) -----------------------------------------------
 * The reference's step contract (PokerGPU.py:527-633: int64 actions in; float32 [N, 13 + 3 (M - 1)] observations,
 * float32 rewards, bool dones out) for callers whose arrays live in host memory.  Per table-step 1 B goes to the
 * device (the action id) and, only for tables that were live when the step began, one packed record comes back
 * (csrc/poker_wire.cuh: 28 / 52 / 76 B for 2 / 6 / 10 seats, compacted in table order); a pool of host threads
 * expands the records into the caller's arrays.  A table that was already finished keeps its row in obs_host (the
 * reference returns the same row for it), reward 0, done 1 -- so the SAME obs / reward / done arrays must be passed to
 * every call of one env (the reference's `self.obs` is one buffer too, :176-179).  The arrays may be pageable.
 * The handle owns device state, pinned staging buffers, streams (one per chunk: copies and kernels of chunk c+1 overlap
 * the expansion of chunk c) and threads.  Unlike the rest of this header these calls allocate (create), block until
 * the outputs are complete (reset / step) and use their own streams.  Whole-batch hands only (no auto_reset).
 * Requires n_players * max_bbs < 65536 (every chip count must fit the record's 16-bit fields) and (n_players, max_players) in
 * {(2,2), (3..4,4), (5..6,6), (7..10,10)}; other shapes use the device-buffer API above.
 * host_hand_ranks: the HandRanks table (pulse_handranks_generate), at least its first 36,875 entries.
 * n_chunks / n_threads < 1: 1 chunk / one thread per online CPU. */
typedef struct PulsePokerHostEnv PulsePokerHostEnv;
int pulse_poker_host_create(PulsePokerHostEnv** out, const PulsePokerConfig* cfg, const int32_t* host_hand_ranks,
                            size_t n_hand_ranks, uint64_t seed, uint64_t env_base, int device, int n_chunks,
                            int n_threads);
void pulse_poker_host_destroy(PulsePokerHostEnv* env);
/* reset (PokerGPU.py:73-157) with decks from the chance tape; active_players = the table size of this hand.
 * obs_host == NULL: records only (every table's record is left in the pinned buffer, see pulse_poker_host_step_records) */
int pulse_poker_host_reset(PulsePokerHostEnv* env, int active_players, int rotation, float* obs_host,
                           float* reward_host_or_null, uint8_t* done_host_or_null);
/* step (:527-633); actions_host: uint8 [N] (action_bytes 1) or int64 [N] (action_bytes 8).  A uint8 array that is
 * page-locked (cudaHostAlloc / cudaHostRegister) is handed to the copy engine where it lies; anything else is staged
 * into the env's own pinned buffer first. */
int pulse_poker_host_step(PulsePokerHostEnv* env, const void* actions_host, int action_bytes, float* obs_host,
                          float* reward_host, uint8_t* done_host);
/* The step without the float32 expansion, for consumers that read the packed records themselves (csrc/poker_wire.cuh; or
 * expand single rows on demand with pulse_poker_wire_expand_host): only done_host is written.  The records of the call
 * stay in the env's pinned buffer until the next call: chunk c's records (table order), their number and the chunk's
 * table range come from pulse_poker_host_records; a table has a record iff its bit is set in pulse_poker_host_live_mask
 * (bit b of word w = table 32 w + b was live when the call began; after a reset every table has one).  Do not mix with
 * pulse_poker_host_step on one env between resets (the float32 arrays would miss the skipped steps). */
int pulse_poker_host_step_records(PulsePokerHostEnv* env, const void* actions_host, int action_bytes, uint8_t* done_host);
int pulse_poker_host_records(const PulsePokerHostEnv* env, int chunk, const uint32_t** rows, long long* n_rows,
                             long long* first_table, long long* n_tables);
const uint32_t* pulse_poker_host_live_mask(const PulsePokerHostEnv* env);
/* record size, observation width, chunk / thread counts, and the bytes the last call moved over PCIe */
int pulse_poker_host_info(const PulsePokerHostEnv* env, int* row_bytes, int* obs_width, int* chunks, int* threads,
                          long long* last_h2d_bytes, long long* last_d2h_bytes);
/* where the time of pulse_poker_host_step goes, accumulated in seconds: out6 = {calls, staging the actions, submitting
 * GPU work, expansion phase (wall clock), worker seconds waiting for records, worker seconds expanding}; reset != 0
 * clears the counters */
int pulse_poker_host_timing(PulsePokerHostEnv* env, double* out6, int reset);
/* diagnostic: per chunk c, seconds (summed over the calls since the last timing reset) after a call's expansion phase
 * began at which the last worker saw the chunk's records land (out[2 c]) and finished its slice (out[2 c + 1]); returns
 * the number of chunks */
int pulse_poker_host_chunk_timing(const PulsePokerHostEnv* env, double* out, int max_chunks);
/* pure host function: n_rows packed records (pkw_row_words(M) 32-bit words each) -> float32 rows [n_rows, 13 + 3 (M - 1)],
 * rewards, dones.  vectorised != 0: through the host env's own expander (AVX2 when the CPU has it, see
 * pulse_poker_host_expand_isa), else the portable definition -- the two must agree bit for bit. */
int pulse_poker_wire_expand_host(const uint32_t* rows, long long n_rows, int max_players, int vectorised,
                                 float* obs_out, float* reward_out, uint8_t* done_out);
/* pure host function: one step of the host env's own bookkeeping without a GPU.  rows: the records of the tables whose
 * bit is set in live_inout, compacted in table order (of every table when fresh != 0); obs / reward / done: the caller's
 * persistent arrays; live_inout / prev_live_inout: ceil(n_tables / 32) words, updated as the env updates them. */
int pulse_poker_wire_expand_step_host(const uint32_t* rows, uint32_t* live_inout, uint32_t* prev_live_inout,
                                      long long n_tables, int max_players, int fresh, float* obs, float* reward,
                                      uint8_t* done);
const char* pulse_poker_host_expand_isa(void);

/* ---- host-buffer 2048 (the end-to-end path of the batched 2048 env) ------------------------------------------------
 * TFE.reset / TFE.step (/root/reference/environments/2048/TFE.py:143-189) for N boards whose arrays live in HOST memory
 * (pageable is fine).  What crosses PCIe is 1 B of action per board in and the engine's packed board (8 B) + reward
 * (4 B) + done (1 B) out; host threads rebuild the reference's int32 tile values [N, side, side] (:121) in the
 * caller's array, chunk by chunk while the next chunk's step and copies run.  Chance from the tape only, keyed by
 * (seed, env_base + board, call counter): results equal pulse_tfe_reset / pulse_tfe_step called once per reset / step
 * with counters 0, 1, 2, ...  auto_reset as pulse_tfe_step.  n_chunks / n_threads: 0 = one chunk / all host threads.
 * The handle owns device state, pinned staging, streams and threads; calls on one handle are serial and synchronous. */
typedef struct PulseTfeHostEnv PulseTfeHostEnv;
int pulse_tfe_host_create(PulseTfeHostEnv** out, long long n_envs, int side, int auto_reset, uint64_t seed,
                          uint64_t env_base, int device, int n_chunks, int n_threads);
void pulse_tfe_host_destroy(PulseTfeHostEnv* env);
/* obs_host int32 [N, side, side] or NULL; done_host_or_null uint8 [N] (cleared) */
int pulse_tfe_host_reset(PulseTfeHostEnv* env, int32_t* obs_host, uint8_t* done_host_or_null);
/* actions_host uint8 or int64 [N] (action_bytes 1 or 8); obs_host int32 [N, side, side] or NULL (boards only, see
 * pulse_tfe_host_boards); reward_host int32 [N]; done_host uint8 [N] */
int pulse_tfe_host_step(PulseTfeHostEnv* env, const void* actions_host, int action_bytes, int32_t* obs_host,
                        int32_t* reward_host, uint8_t* done_host);
const uint64_t* pulse_tfe_host_boards(const PulseTfeHostEnv* env);   /* packed boards of the last call, pinned host memory */
int pulse_tfe_host_scores(PulseTfeHostEnv* env, int64_t* score_host); /* TFE.total_score (:183) of every board */
int pulse_tfe_host_info(const PulseTfeHostEnv* env, int* chunks, int* threads, long long* last_h2d_bytes,
                        long long* last_d2h_bytes);
/* pure host function: n packed boards -> int32 tile values [n, side, side], through the env's own expander */
int pulse_tfe_expand_host(const uint64_t* boards, long long n, int side, int32_t* obs_out);

/* ---- run statistics (the vector that is all-reduced across GPUs) -------------------
 * stats4 double [4] on the device: += {n, number of done flags set, sum(reward), sum(reward^2)}.
 * reward: float32 [n] (reward_is_float != 0) or int32 [n]. */
int pulse_stats_accumulate(const void* reward, int reward_is_float, const uint8_t* done_or_null, double* stats4,
                           long long n, pulse_stream_t stream);

/* ---- hand-level statistics ----------------------------------------------------
 * One pass per step over the hands that finished on this step (done && !terminated), replacing the
 * host-side mask gathers of /root/reference/scripts/Poker/trainGPU_performance.py:186-206 and the reductions
 * of /root/reference/utils/performance.py:55-59, :104-253.  stats67 double [67] on the device, accumulated:
 *   [0] hands  [1] sum d  [2] sum d^2  [3] #(d > 0)        d = stacks[q_seat] - initial_q_stacks
 *   per terminal-stage bucket b (stage >= 4 -> 4): [4+b] hands  [9+b] sum d  [14+b] wins
 *   per position p = (q_seat - button) mod active_players: [19+p] hands  [35+p] sum d  [51+p] wins
 * terminated_inout uint8 [N] |= done (the caller's `terminated`).  The vector is what gets all-reduced.
 * rearm != 0 (packed engine with auto_reset): every raised done flag is a finished hand, `terminated` is not
 * used, and initial_q_stacks[g] is set to the stack the next deal will start from (refilled like
 * PokerGPU.py:105-108), so consecutive hands of a table are each measured from their own start. */
#define PULSE_HAND_STATS_LEN 67
int pulse_poker_hand_stats(const int32_t* stacks /*[N,P]*/, int n_players, const int32_t* stages /*[N]*/,
                           const int32_t* button /*[N]*/, const uint8_t* done, uint8_t* terminated_inout,
                           int32_t* initial_q_stacks /*[N]*/, int q_seat, int active_players, double* stats67,
                           long long n_games, pulse_stream_t stream);
int pulse_poker_packed_hand_stats(const PulsePokerPacked* st, const PulsePokerConfig* cfg, const uint8_t* done,
                                  uint8_t* terminated_inout, int32_t* initial_q_stacks, int q_seat, int rearm,
                                  double* stats67, pulse_stream_t stream);

/* ---- transition hand-off to the learner ------------------------------------------
 * Appends this step's transitions of the learning seat -- tables with seat_idx_before == q_seat that were not
 * terminated before the step -- to a device ring buffer, in table order: the masked gathers of
 * /root/reference/scripts/Poker/trainGPU.py:79-94 plus ReplayBufferTorch.add
 * (/root/reference/utils/ReplayBuffer.py:47-60: slots arange(position, position + batch) % capacity).
 * state / next_state float32 [N, obs_width] (the observation before and after the step), actions int64 [N],
 * rewards float32 [N], dones uint8 [N], seat_idx_before int32 [N], terminated_before uint8 [N] or NULL.
 * Ring: state_buf / next_state_buf float32 [capacity, obs_width], action_buf int64, reward_buf float32,
 * done_buf uint8 [capacity].  total_inout uint64 [1] on the device = transitions appended so far
 * (position = total % capacity, size = min(total, capacity)); appended_out uint64 [1] = rows added by this call.
 * scratch: pulse_replay_scratch_bytes(n_games) bytes of device memory.  Requires n_games <= capacity. */
size_t pulse_replay_scratch_bytes(long long n_games);
int pulse_replay_append(const float* state, const float* next_state, int obs_width, const int64_t* actions,
                        const float* rewards, const uint8_t* dones, const int32_t* seat_idx_before,
                        const uint8_t* terminated_before_or_null, int q_seat, float* state_buf, float* next_state_buf,
                        int64_t* action_buf, float* reward_buf, uint8_t* done_buf, long long capacity,
                        uint64_t* total_inout, uint64_t* appended_out, void* scratch, long long n_games,
                        pulse_stream_t stream);
/* pulse_replay_append with no allocation anywhere -- no counterpart in the reference; added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments and results as pulse_replay_append minus `scratch`: the scratch is a
 * file-scope device array of the library (65,792 B per GPU that loads the library, held until exit), so the entry
 * point calls no cudaMalloc / cudaFree / allocating library routine.  Limit: n_games <= pulse_replay_static_max_games()
 * = 4,194,304 tables per call; larger batches are refused with PULSE_EINVAL and a message in pulse_last_error().
 * All calls on one device share the array and nothing in it needs clearing between calls, but two calls that overlap
 * on the GPU (different streams without a dependency) race: keep them on one stream or order them with events. */
long long pulse_replay_static_max_games(void);
int pulse_replay_append_static(const float* state, const float* next_state, int obs_width, const int64_t* actions,
                               const float* rewards, const uint8_t* dones, const int32_t* seat_idx_before,
                               const uint8_t* terminated_before_or_null, int q_seat, float* state_buf,
                               float* next_state_buf, int64_t* action_buf, float* reward_buf, uint8_t* done_buf,
                               long long capacity, uint64_t* total_inout, uint64_t* appended_out, long long n_games,
                               pulse_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif

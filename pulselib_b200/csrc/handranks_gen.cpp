// Regenerates the TwoPlusTwo seven-card evaluator table ("HandRanks.dat", 32,487,834 int32) that
// the reference loads from an un-versioned, git-ignored download
// (/root/reference/environments/Poker/PokerGPU.py:47-58; .gitignore:6).  This is a restatement of
// the published generator algorithm (Ray Wotton's generate_table: 64-bit card-set ids, one byte per
// card = ((rank+1) << 4) | suit, suits zeroed once they can no longer make a flush, ids kept in
// numeric order, 53 int32 slots per id), written breadth-first instead of insertion-sorted:
//   HR[53 + node*53 + card] = 53 + next_node*53   for ids of fewer than 7 cards
//                           = hand value           for the 7th card
//   HR[53 + node*53 + 0]    = hand value of the id itself for 5- and 6-card ids
// Layout matters only to the reference's flop "equity" HR[HR[p5]] (PokerGPU.py:521), which
// re-uses a hand value as an index into the low (1-3 card) part of the table.
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <vector>

#include "common.cuh"
#include "poker_eval.cuh"

namespace {

constexpr size_t kEntries = 32487834;

// id of (id_in + newcard), 0 if the card is already present or would be a fifth of its rank.
// *n_out = number of cards in the new id.
uint64_t make_id(uint64_t id_in, int newcard, int* n_out) {
  int work[8] = {0};
  int suitcount[5] = {0}, rankcount[14] = {0};
  for (int i = 0; i < 6; ++i) work[i + 1] = (int)((id_in >> (8 * i)) & 0xff);
  int c = newcard - 1;
  work[0] = (((c >> 2) + 1) << 4) + (c & 3) + 1;
  int n = 0;
  bool dup = false;
  for (n = 0; work[n]; ++n) {
    suitcount[work[n] & 0xf]++;
    rankcount[(work[n] >> 4) & 0xf]++;
    if (n && work[0] == work[n]) dup = true;
  }
  *n_out = n;
  if (dup) return 0;
  int needsuited = n - 2;
  if (n > 4)
    for (int r = 1; r < 14; ++r)
      if (rankcount[r] > 4) return 0;
  if (needsuited > 1)
    for (int i = 0; i < n; ++i)
      if (suitcount[work[i] & 0xf] < needsuited) work[i] &= 0xf0;
  std::sort(work, work + 7, [](int a, int b) { return a > b; });
  uint64_t id = 0;
  for (int i = 0; i < 7; ++i) id |= (uint64_t)work[i] << (8 * i);
  return id;
}

// Hand value of a 5/6/7-card id.  Cards whose suit byte was zeroed cannot take part in a flush.
int eval_id(uint64_t id) {
  if (!id) return 0;
  int cnt[13] = {0};
  uint32_t suitmask[5] = {0};
  int n = 0;
  for (int i = 0; i < 7; ++i) {
    int b = (int)((id >> (8 * i)) & 0xff);
    if (!b) break;
    int r = (b >> 4) - 1;
    cnt[r]++;
    suitmask[b & 0xf] |= 1u << r;
    ++n;
  }
  uint32_t any = 0, pairs = 0, trips = 0, quads = 0, flush = 0;
  for (int r = 0; r < 13; ++r) {
    if (cnt[r]) any |= 1u << r;
    if (cnt[r] == 2) pairs |= 1u << r;
    if (cnt[r] == 3) trips |= 1u << r;
    if (cnt[r] == 4) quads |= 1u << r;
  }
  for (int s = 1; s <= 4; ++s)
    if (pe_popc(suitmask[s]) >= 5) flush = suitmask[s];
  return pe_eval_core(any, pairs, trips, quads, flush);
}

}  // namespace

extern "C" {

size_t pulse_handranks_entries(void) { return kEntries; }

// Writes the table into out[0..n_entries-1]; n_entries must be pulse_handranks_entries().
// Host-only, single call ~ a few seconds.
int pulse_handranks_generate(int32_t* out, size_t n_entries) {
  PULSE_REQUIRE(out != nullptr, "out is null");
  PULSE_REQUIRE(n_entries == kEntries, "n_entries must equal pulse_handranks_entries()");
  std::vector<std::vector<uint64_t>> level(7);
  level[0].push_back(0);
  int n;
  for (int l = 0; l < 6; ++l) {
    std::vector<uint64_t>& next = level[l + 1];
    next.reserve(level[l].size() * 8);
    for (uint64_t id : level[l])
      for (int card = 1; card <= 52; ++card) {
        uint64_t nid = make_id(id, card, &n);
        if (nid) next.push_back(nid);
      }
    std::sort(next.begin(), next.end());
    next.erase(std::unique(next.begin(), next.end()), next.end());
  }
  std::vector<uint64_t> ids;
  size_t base[8] = {0};
  for (int l = 0; l <= 6; ++l) {
    base[l] = ids.size();
    ids.insert(ids.end(), level[l].begin(), level[l].end());
  }
  base[7] = ids.size();
  if (ids.size() * 53 + 53 > kEntries + 52) {
    pulse_set_error("pulse_handranks_generate: %zu ids do not fit the table", ids.size());
    return PULSE_EINVAL;
  }
  memset(out, 0, kEntries * sizeof(int32_t));
  for (int l = 0; l <= 6; ++l) {
    const std::vector<uint64_t>& nxt = level[l < 6 ? l + 1 : 6];
    for (size_t i = base[l]; i < base[l + 1]; ++i) {
      uint64_t id = ids[i];
      for (int card = 1; card <= 52; ++card) {
        uint64_t nid = make_id(id, card, &n);
        int32_t slot;
        if (l < 6) {
          if (!nid) slot = 53;  // node 0 * 53 + 53, as the original's SaveID(0) == 0
          else {
            size_t pos = std::lower_bound(nxt.begin(), nxt.end(), nid) - nxt.begin();
            slot = (int32_t)((base[l + 1] + pos) * 53 + 53);
          }
        } else {
          slot = eval_id(nid);
        }
        size_t at = i * 53 + card + 53;
        if (at < kEntries) out[at] = slot;
      }
      if (l == 5 || l == 6) out[i * 53 + 53] = eval_id(id);
    }
  }
  return PULSE_OK;
}

// Walks the table exactly as the reference does (p = 53; p = HR[p + card] ...), for n_cards in
// 5..7; host helper for tools and tests.
int pulse_handranks_walk(const int32_t* hr, const int32_t* cards, int n_cards) {
  int p = 53;
  for (int i = 0; i < n_cards; ++i) p = hr[p + cards[i]];
  if (n_cards < 7) p = hr[p];
  return p;
}

// Computed evaluator on the host (same code the kernels run): cards 1..52, n_cards in 5..7.
int pulse_eval_cards(const int32_t* cards, int n_cards) {
  uint32_t m[4] = {0, 0, 0, 0};
  for (int i = 0; i < n_cards; ++i) pe_add_card(cards[i], m);
  return pe_eval_suits(m[0], m[1], m[2], m[3]);
}
}

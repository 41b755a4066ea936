// Units of a launch and the work list of the fused kernel (host side; plain C++, shared with the CPU tests).
#pragma once
#include <stdint.h>
#include <algorithm>
#include <vector>

namespace mmb {

// One unit of independent work: a track (batch mode) or a frame range of a long signal (shard mode).
// Mirrors mmb_track in include/madmom_b200.h.
struct Track {
  int64_t pcm_offset;    // element offset of local sample 0 inside the PCM buffer
  int64_t pcm_len;       // number of local samples; anything outside reads as zero (signal.py:928-962)
  int64_t sample_base;   // global sample index of local sample 0 (0 for whole tracks)
  int64_t frame_first;   // global index of the first output frame
  int64_t frame_count;   // number of output rows
  int64_t out_row;       // first output row of this unit
};

// A run of consecutive output frames of one unit, walked by one warp group in rounds.  The host cuts the units of a
// launch into chunks and lays them out so that the groups get equal numbers of rounds (build_work_list below); an
// entry with no frames is a hole in that lay-out and is skipped.  Results do not depend on the cuts.
struct Chunk {
  int32_t unit;          // index into the unit table
  int32_t frames;        // output frames; 0 -> hole
  int64_t first;         // first output frame, local to the unit (0 .. frame_count)
};

// Work list of the fused kernel for one launch.  Warp group v of the V = grid * groups resident ones walks the
// entries v, v + V, ... in rounds of `fpr` frames, and the time per round is the same everywhere (6.9 us for the
// SuperFlux recipe, to 1 % whatever the cuts), so the launch takes as long as the group with the most rounds.  Equal
// chunks dealt in turn leave that group 1 .. 18 % above the mean (1024 x 3000 frames: 1316 rounds against 1301; 64
// tracks: 96 against 82; tracks of mixed lengths: 7 .. 9 %).  So the list is laid out per group instead: every group
// gets the same number of rounds, as one run of consecutive frames cut at the unit boundaries.  A cut inside a unit
// costs its lead-in frames (`lead` = diff_frames) a second time, which the budget counts.  A launch with less than
// one round per group uses fewer CTAs (*grid).
// fixed_chunk > 0: equal chunks of that many frames in unit order instead (mmb_plan_config::chunk_frames).
inline void build_work_list(int N, int fpr, int groups, int num_sms, int64_t lead, int64_t fixed_chunk,
                            int64_t wave_rounds, const Track* u, int32_t n, std::vector<Chunk>& out, int* grid) {
  const int64_t slots = (int64_t)num_sms * groups;
  out.clear();
  if (fixed_chunk > 0) {
    const int64_t c = fixed_chunk;
    for (int32_t i = 0; i < n; ++i)
      for (int64_t a = 0; a < u[i].frame_count; a += c)
        out.push_back(Chunk{i, (int32_t)std::min<int64_t>(c, u[i].frame_count - a), a});
    *grid = (int)std::min<int64_t>(num_sms, ((int64_t)out.size() + groups - 1) / groups);
    return;
  }
  auto rounds_of = [&](int64_t frames) { return (frames + lead + fpr - 1) / fpr; };
  // a run of frames is cut into entries of at most kMaxFrames (the entry holds an int32)
  const int64_t kMaxFrames = (int64_t)1 << 30;
  std::vector<std::vector<Chunk>> list(slots);
  std::vector<int64_t> after(n + 1, 0);                     // rounds of the units behind unit i
  for (int32_t i = n - 1; i >= 0; --i) after[i] = after[i + 1] + (u[i].frame_count > 0 ? rounds_of(u[i].frame_count) : 0);
  int32_t i = 0;
  int64_t a = 0;                                            // next frame: unit i, local frame a
  auto skip_empty = [&]() { while (i < n && u[i].frame_count == 0) ++i; };
  skip_empty();
  // one wave: every group takes `budget` rounds from the front of what is left
  auto wave = [&](int64_t budget) {
    for (int64_t s = 0; s < slots && i < n; ++s) {
      int64_t b = budget;
      while (b > 0 && i < n) {
        const int64_t rest = u[i].frame_count - a, need = rounds_of(rest);
        int64_t take = rest;
        if (need > b) {
          take = b * fpr - lead;
          // N = 1024 pairs frames by their global index: cut in front of an even one (next_round)
          if (N == 1024 && ((u[i].frame_first + a + take) & 1)) --take;
          if (take <= 0) break;
          b = 0;
        } else {
          b -= need;
        }
        for (int64_t done = 0; done < take; done += kMaxFrames)
          list[s].push_back(Chunk{i, (int32_t)std::min(kMaxFrames, take - done), a + done});
        a += take;
        if (a >= u[i].frame_count) { a = 0; ++i; skip_empty(); }
      }
    }
  };
  const int64_t min_budget = lead / fpr + 1;
  std::vector<size_t> mark(slots);
  while (i < n) {
    const int64_t left = rounds_of(u[i].frame_count - a) + after[i + 1];
    const int64_t waves = wave_rounds > 0 ? std::max<int64_t>((left + slots * wave_rounds / 2) / (slots * wave_rounds), 1) : 1;
    int64_t budget = std::max((left + slots * waves - 1) / (slots * waves), min_budget);
    if (waves > 1) {
      const int32_t i_was = i;
      const int64_t a_was = a;
      wave(budget);
      if (i != i_was || a != a_was) continue;               // (no progress: lead-in longer than the budget; finish below)
    }
    // the last wave takes all that is left; its cuts add a round here and there: one or two more tries
    const int32_t i0 = i;
    const int64_t a0 = a;
    for (int64_t s = 0; s < slots; ++s) mark[s] = list[s].size();
    for (;; ++budget) {
      wave(budget);
      if (i >= n) break;
      i = i0; a = a0;
      for (int64_t s = 0; s < slots; ++s) list[s].resize(mark[s]);
    }
  }
  int64_t used = 0;
  size_t depth = 0;
  for (int64_t s = 0; s < slots; ++s)
    if (!list[s].empty()) { used = s + 1; depth = std::max(depth, list[s].size()); }
  *grid = (int)((used + groups - 1) / groups);
  const int64_t stride = (int64_t)*grid * groups;
  out.assign(depth * stride, Chunk{0, 0, 0});
  for (int64_t s = 0; s < used; ++s)
    for (size_t k = 0; k < list[s].size(); ++k) out[k * stride + s] = list[s][k];
}

}  // namespace mmb

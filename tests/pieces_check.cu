// host-only check of the piece arithmetic: every block of both segments covered exactly once,
// coarse pieces full, slots in row order
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>
#include "../dstack_b200/csrc/kernels.cuh"
int main() {
  srand(7);
  for (int it = 0; it < 20000; ++it) {
    Pieces pc{};
    pc.n_groups = 1 + rand() % 200;
    uint32_t a_blocks = rand() % 3 == 0 ? 0 : rand() % 70000, b_blocks = rand() % 4 == 0 ? 0 : rand() % 20000;
    if (a_blocks + b_blocks == 0) a_blocks = 1;
    pc.n_blocks = a_blocks + b_blocks; pc.split_block = a_blocks;
    pc.wpw = 1 + rand() % 40; pc.wpw_f = 1 + rand() % pc.wpw;
    uint32_t wins_b = Pieces::windows(b_blocks), wins_a = Pieces::windows(a_blocks);
    uint32_t fine_total = rand() % 5 == 0 ? 0 : rand() % (wins_a + wins_b + 5);
    uint32_t fine_b = std::min(wins_b, fine_total), fine_a = fine_total - fine_b;
    pc.pa = pc.cut(a_blocks, fine_a, pc.ca);
    pc.pb = pc.cut(b_blocks, fine_b, pc.cb);
    std::vector<int> cover(pc.n_blocks, 0);
    uint32_t prev_end = 0;
    for (uint32_t pos = 0; pos < pc.per_group(); ++pos)
      for (uint32_t w = 0; w < MATCH_WARPS; ++w) {
        uint32_t b0, b1; pc.warp_blocks(pos, w, b0, b1);
        if (b0 > b1) { printf("neg range\n"); return 1; }
        if (b0 < b1 && b0 != prev_end) { printf("gap/overlap it %d pos %u w %u: %u vs %u\n", it, pos, w, b0, prev_end); return 1; }
        if (b0 < b1) prev_end = b1;
        if ((b1 - b0 + WIN_BLOCKS - 1) / WIN_BLOCKS > pc.wpw) { printf("too many windows\n"); return 1; }
        if (b0 < b1 && (b0 < a_blocks) != (b1 <= a_blocks)) { printf("straddles the split\n"); return 1; }
        if (b0 < b1 && (pos < pc.pa) != (b0 < a_blocks)) { printf("segment mismatch\n"); return 1; }
        for (uint32_t b = b0; b < b1; ++b) cover[b]++;
      }
    for (uint32_t b = 0; b < pc.n_blocks; ++b) if (cover[b] != 1) { printf("block %u covered %d times (it %d)\n", b, cover[b], it); return 1; }
  }
  printf("pieces ok\n");
  return 0;
}

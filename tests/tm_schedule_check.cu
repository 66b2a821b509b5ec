// Host-only check of the decoder kernel's schedule tables (tm_build_tables, csrc/ldpc_kernels_tmem.cuh) for every base graph,
// lifting size and TMEM budget: compiled and run by tests/test_tm_schedule.py (nvcc, no GPU). Prints "ok <n>" or the first
// violation.
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <map>
#include <set>
#include <utility>
#include <vector>
#define LDPC_TM_DEVICE_ONLY
#include "ldpc_kernels_tmem.cuh"

using namespace ldpc_b200;

#define CHECK(cond, ...)                       \
  do {                                         \
    if (!(cond)) {                             \
      std::printf("FAIL bg=%d Z=%d cap=%u: ", bg, Z, cap); \
      std::printf(__VA_ARGS__);                \
      std::printf("\n");                       \
      return 1;                                \
    }                                          \
  } while (0)

static int check(int bg, int Z, unsigned cap)
{
  const int M = bg == 2 ? LDPC_BG2_M : LDPC_BG1_M, K = bg == 2 ? 10 : 22, hrr = K + 4;
  const uint16_t* row_ptr = bg == 2 ? LDPC_BG2_ROW_PTR : LDPC_BG1_ROW_PTR;
  const uint8_t* col = bg == 2 ? LDPC_BG2_COL : LDPC_BG1_COL;
  int ls = -1;
  for (int i = 0; i != LDPC_NOF_LIFTING_SIZES; ++i) {
    if (LDPC_LIFTING_SIZES[i] == Z) {
      ls = LDPC_LIFTING_SET[i];
    }
  }
  const uint16_t* shift = bg == 2 ? LDPC_BG2_SHIFT[ls] : LDPC_BG1_SHIFT[ls];
  TmemParams p;
  std::memset(&p, 0, sizeof(p));
  CHECK(tm_build_tables(bg, Z, ls, p, cap), "tm_build_tables refused the graph (the launch would fall back to the general kernel)");
  // rows: class, edges (the reference's adjacency as a multiset), message words
  std::map<unsigned, int> row_of_edge0;
  std::vector<std::set<unsigned>> cols(M);
  std::vector<std::pair<unsigned, unsigned>> words;
  for (int r = 0; r != M; ++r) {
    const unsigned cls = p.rowdesc[r] & 0xffu, e0 = (p.rowdesc[r] >> 8) & 0xfffu, tmw = p.rowdesc[r] >> 20, d = cls >> 1;
    std::multiset<std::pair<unsigned, unsigned>> want, got;
    unsigned ext = 0;
    for (int e = row_ptr[r]; e != row_ptr[r + 1]; ++e) {
      if (col[e] < hrr) {
        want.insert({col[e], shift[e] % (unsigned)Z});
      } else {
        ++ext;
      }
    }
    CHECK(ext == (cls & 1u) && want.size() == d, "row %d: class %u against %zu high-rate edges, %u extension", r, cls, want.size(), ext);
    for (unsigned k = 0; k != d; ++k) {
      const TmEdge ed = p.edge[e0 + k];
      got.insert({ed.cb / (4u * (unsigned)Z), ed.s4 / 4u});
      cols[r].insert(ed.cb / (4u * (unsigned)Z));
    }
    CHECK(got == want, "row %d: edge records differ from the base graph", r);
    row_of_edge0[e0] = r;
    words.push_back({tmw, tmw + (d + 1) / 2});
  }
  for (size_t i = 0; i != words.size(); ++i) {
    for (size_t j = i + 1; j != words.size(); ++j) {
      CHECK(words[i].second <= words[j].first || words[j].second <= words[i].first, "message words of rows %zu and %zu overlap", i, j);
    }
    const bool in_tmem = words[i].second <= (cap & ~7u), in_ovf = words[i].first >= cap;
    CHECK(in_tmem || in_ovf, "row %zu: message words straddle the TMEM budget", i);
    CHECK(!in_ovf || words[i].second - cap <= p.ovf_words, "row %zu: overflow words beyond ovf_words", i);
  }
  // steps: every row once, in order; pairs adjacent, orthogonal and of a compiled shape
  int next_row = 0;
  unsigned s = 0;
  std::vector<int> rows_after;  // rows covered after step s
  for (; s <= (unsigned)M && p.steps[s].x != 0; ++s) {
    const uint4 st = p.steps[s];
    const unsigned cls = st.x & 0xffu, nrows = (st.x >> 8) & 0xffu;
    CHECK(nrows == 1 || nrows == 2, "step %u: %u rows", s, nrows);
    CHECK(row_of_edge0.count(st.y & 0xffffu), "step %u: row A's first edge is no row's", s);
    const int rA = row_of_edge0[st.y & 0xffffu];
    CHECK((st.y >> 16) == (p.rowdesc[rA] >> 20), "step %u: message word of row A", s);
    const unsigned zpad4 = 4u * (((unsigned)Z + 31u) / 32u) * 32u;
    CHECK((st.w & 0xffffu) == (rA >= 4 ? (unsigned)(rA - 4) * zpad4 : 0u), "step %u: extension offset of row A", s);
    if (nrows == 1) {
      CHECK(rA == next_row && cls == (p.rowdesc[rA] & 0xffu), "step %u: single row %d, expected %d", s, rA, next_row);
      CHECK(ldpc_b200::tm::tm_row_used(bg, (int)(cls >> 1), (cls & 1u) != 0), "step %u: row shape %u not compiled for this base graph", s, cls);
    } else {
      CHECK(row_of_edge0.count(st.z & 0xffffu), "step %u: row B's first edge is no row's", s);
      const int rB = row_of_edge0[st.z & 0xffffu];
      const int dA = (int)((p.rowdesc[rA] & 0xffu) >> 1), dB = (int)((p.rowdesc[rB] & 0xffu) >> 1);
      CHECK(std::min(rA, rB) == next_row && std::max(rA, rB) == next_row + 1 && next_row >= 4, "step %u: pair (%d, %d), expected %d", s, rA, rB, next_row);
      CHECK(dA >= dB && cls == (unsigned)(16 * dA + dB) && ldpc_b200::tm::tm_pair_used(bg, dA, dB), "step %u: pair class %u for degrees %d, %d", s, cls, dA, dB);
      CHECK((st.z >> 16) == (p.rowdesc[rB] >> 20) && (st.w >> 16) == (unsigned)(rB - 4) * zpad4, "step %u: row B's fields", s);
      for (unsigned c : cols[rA]) {
        CHECK(!cols[rB].count(c), "step %u: rows %d and %d share column %u", s, rA, rB, c);
      }
    }
    next_row += (int)nrows;
    rows_after.push_back(next_row);
  }
  CHECK(next_row == M, "steps cover %d of %d rows", next_row, M);
  CHECK(p.steps[s].x == 0 && p.steps[s].y == 0, "no sentinel behind the last step");
  // layer limits: the steps that cover rows [0, L), whole steps only
  for (int L = 1; L <= M; ++L) {
    const unsigned n = p.layer_steps[L] & 0xffu, covered = p.layer_steps[L] >> 8;
    CHECK(n >= 1 && n <= s && (int)covered == rows_after[n - 1], "layer_steps[%d] = %u steps / %u rows", L, n, covered);
    CHECK((int)covered >= L && (n == 1 || rows_after[n - 2] < L), "layer_steps[%d]: %u rows in %u steps is not the shortest cover", L, covered, n);
  }
  return 0;
}

int main()
{
  int n = 0;
  for (int bg = 1; bg <= 2; ++bg) {
    for (int i = 0; i != LDPC_NOF_LIFTING_SIZES; ++i) {
      for (unsigned cap : {(unsigned)TM_COLS_PER_WARP, (unsigned)TM_COLS_PER_WARP_WIDE}) {
        if (check(bg, LDPC_LIFTING_SIZES[i], cap) != 0) {
          return 1;
        }
        ++n;
      }
    }
  }
  std::printf("ok %d\n", n);
  return 0;
}

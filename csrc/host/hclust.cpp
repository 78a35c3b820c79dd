// Agglomerative clustering by the nearest-neighbour-chain algorithm on a precomputed distance matrix (work row 28).
//
// The reference orders its heatmaps with ComplexHeatmap's `cluster_rows / cluster_columns = TRUE`
// (composite_code/rnotebook/cml_targetaml_seanalysis.Rmd:536-613, 927-945), i.e. hclust(dist(x), "complete").  Here the
// distance matrix comes from the tcgen05 squared-distance GEMM (ops/gemm.py sq_dists); the merge sequence itself is
// inherently sequential (every merge changes the distances the next one looks at), O(n^2) work on an n x n matrix that
// is already in memory, so it runs on the host: one chain walk with Lance-Williams updates.
//
// Output convention = scipy.cluster.hierarchy.linkage (the test oracle): merges sorted by height (stable), cluster ids
// n, n+1, ... in that order, the smaller id first in each row; `leaves` is the left-to-right leaf order of the dendrogram.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <limits>
#include <numeric>
#include <vector>

namespace {

enum Method { SINGLE = 0, COMPLETE = 1, AVERAGE = 2, WEIGHTED = 3, WARD = 4 };

inline double lance_williams(int method, double dxi, double dyi, double dxy, double nx, double ny, double ni) {
  switch (method) {
    case SINGLE: return std::min(dxi, dyi);
    case COMPLETE: return std::max(dxi, dyi);
    case AVERAGE: return (nx * dxi + ny * dyi) / (nx + ny);
    case WEIGHTED: return 0.5 * (dxi + dyi);
    default: {  // Ward on plain (not squared) distances
      const double t = 1.0 / (nx + ny + ni);
      return std::sqrt((ni + nx) * t * dxi * dxi + (ni + ny) * t * dyi * dyi - ni * t * dxy * dxy);
    }
  }
}

struct Merge {
  int a, b;
  double h;
};

}  // namespace

extern "C" {

// dist: n x n row-major, symmetric, zero diagonal (read only).  Z: (n-1) x 4 doubles.  leaves: n ints (may be null).
// Returns 0, or -1 for bad arguments.
int cml_host_hclust(const double* dist, int n, int method, double* Z, int* leaves) {
  if (n < 1 || method < 0 || method > 4 || !dist || (n > 1 && !Z)) return -1;
  if (n == 1) {
    if (leaves) leaves[0] = 0;
    return 0;
  }
  std::vector<double> D(dist, dist + (size_t)n * n);
  std::vector<int> size(n, 1), chain;
  chain.reserve(n);
  std::vector<Merge> merges;
  merges.reserve(n - 1);
  auto d = [&](int i, int j) -> double& { return D[(size_t)i * n + j]; };
  for (int k = 0; k < n - 1; ++k) {
    if (chain.empty()) {
      for (int i = 0; i < n; ++i)
        if (size[i] > 0) { chain.push_back(i); break; }
    }
    int x, y = -1;
    double cur;
    while (true) {
      x = chain.back();
      if (chain.size() > 1) { y = chain[chain.size() - 2]; cur = d(x, y); }      // the previous link wins ties: the chain must close
      else cur = std::numeric_limits<double>::infinity();
      for (int i = 0; i < n; ++i) {
        if (size[i] == 0 || i == x) continue;
        const double v = d(x, i);
        if (v < cur) { cur = v; y = i; }
      }
      if (chain.size() > 1 && y == chain[chain.size() - 2]) break;
      chain.push_back(y);
    }
    chain.pop_back();
    chain.pop_back();
    if (x > y) std::swap(x, y);
    const double nx = size[x], ny = size[y];
    merges.push_back({x, y, cur});
    size[x] = 0;
    size[y] = (int)(nx + ny);
    for (int i = 0; i < n; ++i) {
      if (size[i] == 0 || i == y) continue;
      const double v = lance_williams(method, d(i, x), d(i, y), cur, nx, ny, (double)size[i]);
      d(i, y) = v;
      d(y, i) = v;
    }
  }
  // sort by height (stable), relabel with a union-find exactly like scipy's `label`
  std::vector<int> order(n - 1);
  std::iota(order.begin(), order.end(), 0);
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return merges[a].h < merges[b].h; });
  std::vector<int> parent(2 * n - 1), csize(2 * n - 1, 1);
  std::iota(parent.begin(), parent.end(), 0);
  auto find = [&](int v) {
    int r = v;
    while (parent[r] != r) r = parent[r];
    while (parent[v] != r) { const int nxt = parent[v]; parent[v] = r; v = nxt; }
    return r;
  };
  std::vector<int> left(n - 1), right(n - 1);
  for (int i = 0; i < n - 1; ++i) {
    const Merge& m = merges[order[i]];
    int ra = find(m.a), rb = find(m.b);
    if (ra > rb) std::swap(ra, rb);
    const int id = n + i;
    parent[ra] = id;
    parent[rb] = id;
    csize[id] = csize[ra] + csize[rb];
    left[i] = ra;
    right[i] = rb;
    Z[4 * (size_t)i + 0] = ra;
    Z[4 * (size_t)i + 1] = rb;
    Z[4 * (size_t)i + 2] = m.h;
    Z[4 * (size_t)i + 3] = csize[id];
  }
  if (leaves) {
    // iterative pre-order walk from the root: left subtree first
    std::vector<int> stack;
    stack.push_back(2 * n - 2);
    int pos = 0;
    while (!stack.empty()) {
      const int v = stack.back();
      stack.pop_back();
      if (v < n) { leaves[pos++] = v; continue; }
      stack.push_back(right[v - n]);
      stack.push_back(left[v - n]);
    }
  }
  return 0;
}

}  // extern "C"

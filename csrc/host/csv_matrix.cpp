// Parallel reader for a genes x samples integer CSV (first column = gene id, header = sample ids): the native core of
// io.read_counts_csv.
//
// The reference loads its 60,488 x 474 count matrix with read.csv / pandas (JSmith_code/Differential_Expression_and_Lasso.Rmd:91,
// scripts/archives/david_clean.ipynb cell 20); at that size the single-threaded CSV parse costs seconds - more than every GPU
// stage that follows.  Here the file is mmap-ed, line starts are found with memchr, and a pool of threads parses disjoint row
// ranges straight into the int32 matrix.  Quoted fields ("ENSG...", "12") and "12.0"-style integers are accepted; anything
// else makes the call fail with the line number so that the caller can fall back to a general CSV reader.
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <atomic>
#include <cstdint>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

namespace {

struct Map {
  const char* p = nullptr;
  size_t n = 0;
  int fd = -1;
  bool open(const char* path) {
    fd = ::open(path, O_RDONLY);
    if (fd < 0) return false;
    struct stat st;
    if (fstat(fd, &st) != 0) return false;
    n = (size_t)st.st_size;
    if (n == 0) { p = ""; return true; }
    void* m = mmap(nullptr, n, PROT_READ, MAP_PRIVATE, fd, 0);
    if (m == MAP_FAILED) return false;
    p = static_cast<const char*>(m);
    madvise(m, n, MADV_SEQUENTIAL);
    return true;
  }
  void close_() {
    if (p && n) munmap(const_cast<char*>(p), n);
    if (fd >= 0) close(fd);
    p = nullptr; n = 0; fd = -1;
  }
  ~Map() { close_(); }
};

struct State {
  Map f;
  std::vector<size_t> line;          // start offset of every non-empty line (header first)
  std::vector<std::string> header;   // sample ids (header fields after the id column)
  long rows = 0, cols = 0;
};
State g;

inline void strip(const char*& b, const char*& e) {
  while (b < e && (*b == ' ' || *b == '\t')) ++b;
  while (e > b && (e[-1] == ' ' || e[-1] == '\t' || e[-1] == '\r')) --e;
  if (e - b >= 2 && *b == '"' && e[-1] == '"') { ++b; --e; }
}

inline const char* line_end(const char* p, const char* end) {
  const char* e = static_cast<const char*>(memchr(p, '\n', end - p));
  return e ? e : end;
}

// one data row: id into (idb, ide), `cols` integers into out; returns false on a malformed row
bool parse_row(const char* p, const char* eol, char sep, long cols, int32_t* out, const char*& idb, const char*& ide) {
  const char* fe = static_cast<const char*>(memchr(p, sep, eol - p));
  if (!fe) return false;
  idb = p; ide = fe;
  strip(idb, ide);
  const char* q = fe + 1;
  for (long c = 0; c < cols; ++c) {
    if (q > eol) return false;
    const char* e = (c + 1 < cols) ? static_cast<const char*>(memchr(q, sep, eol - q)) : eol;
    if (!e) return false;
    const char* b = q;
    const char* t = e;
    strip(b, t);
    bool neg = false;
    if (b < t && *b == '-') { neg = true; ++b; }
    long v = 0;
    bool any = false;
    for (; b < t && *b >= '0' && *b <= '9'; ++b) {
      v = v * 10 + (*b - '0');
      any = true;
      if (v > 2147483647L) return false;            // bail out before `v` itself can overflow on a 20-digit field
    }
    if (b < t && *b == '.') {                       // "12.0": digits after the point must all be zero
      ++b;
      for (; b < t && *b == '0'; ++b) {}
    }
    if (!any || b != t || v > 2147483647L) return false;
    out[c] = (int32_t)(neg ? -v : v);
    q = e + 1;
  }
  return true;
}

}  // namespace

extern "C" {

// Step 1: map the file, index its lines, read the header.  Returns 0 and the matrix shape (rows = data lines, cols = header
// fields after the id column), or a negative code: -1 cannot open, -2 empty file / no header.
int cml_host_csv_open(const char* path, char sep, long* rows, long* cols) {
  g.f.close_();
  g.line.clear();
  g.header.clear();
  g.rows = g.cols = 0;
  if (!g.f.open(path)) return -1;
  const char* p = g.f.p;
  const char* end = g.f.p + g.f.n;
  while (p < end) {
    const char* e = line_end(p, end);
    const char* t = e;
    while (t > p && (t[-1] == '\r' || t[-1] == ' ')) --t;
    if (t > p) g.line.push_back((size_t)(p - g.f.p));
    p = e + 1;
  }
  if (g.line.empty()) return -2;
  const char* h = g.f.p + g.line[0];
  const char* he = line_end(h, end);
  bool first = true;
  while (h <= he) {
    const char* fe = static_cast<const char*>(memchr(h, sep, he - h));
    if (!fe) fe = he;
    const char* b = h;
    const char* t = fe;
    strip(b, t);
    if (!first) g.header.emplace_back(b, t);
    first = false;
    if (fe >= he) break;
    h = fe + 1;
  }
  g.rows = (long)g.line.size() - 1;
  g.cols = (long)g.header.size();
  *rows = g.rows;
  *cols = g.cols;
  return 0;
}

// Header field j (sample id) / byte length of the longest row id are read back through these.
int cml_host_csv_header(long j, char* buf, int cap) {
  if (j < 0 || j >= (long)g.header.size()) return -1;
  const std::string& s = g.header[j];
  if ((int)s.size() + 1 > cap) return -2;
  memcpy(buf, s.data(), s.size());
  buf[s.size()] = 0;
  return (int)s.size();
}

// Step 2: parse every data row with n_threads threads.  out: int32 [rows][cols] row-major; ids: rows x id_cap bytes, each
// id NUL-terminated (truncated ids make the call fail).  Returns 0, or -(1-based line number) of the first malformed row.
long cml_host_csv_read(char sep, int32_t* out, char* ids, int id_cap, int n_threads) {
  if (!g.f.p || g.rows <= 0) return -1;
  if (n_threads < 1) n_threads = 1;
  if (n_threads > g.rows) n_threads = (int)g.rows;
  const char* end = g.f.p + g.f.n;
  std::atomic<long> bad{0};
  auto work = [&](long r0, long r1) {
    for (long r = r0; r < r1 && bad.load(std::memory_order_relaxed) == 0; ++r) {
      const char* p = g.f.p + g.line[r + 1];
      const char* eol = line_end(p, end);
      const char* idb;
      const char* ide;
      if (!parse_row(p, eol, sep, g.cols, out + r * g.cols, idb, ide) || (ide - idb) + 1 > id_cap) {
        long expect = 0;
        bad.compare_exchange_strong(expect, r + 2);
        return;
      }
      memcpy(ids + r * (long)id_cap, idb, ide - idb);
      ids[r * (long)id_cap + (ide - idb)] = 0;
    }
  };
  std::vector<std::thread> pool;
  const long per = (g.rows + n_threads - 1) / n_threads;
  for (int t = 0; t < n_threads; ++t) {
    const long r0 = t * per, r1 = r0 + per < g.rows ? r0 + per : g.rows;
    if (r0 < r1) pool.emplace_back(work, r0, r1);
  }
  for (auto& th : pool) th.join();
  return -bad.load();
}

void cml_host_csv_close(void) {
  g.f.close_();
  g.line.clear();
  g.line.shrink_to_fit();
  g.header.clear();
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------- writer
// Matrix -> CSV with a thread pool: every thread formats a block of rows into its own buffer (integers, or doubles in their
// shortest round-trip form via std::to_chars), the blocks are written in order.  The reference writes its 21,407 x 474 TMM
// log2-CPM matrix with write.csv (JSmith_code/Normalization_and_Batch_Effect_Investigation.Rmd:61); pandas needs 21 s for
// that table, this about one.
#include <charconv>
#include <cstdio>

namespace {

template <typename T>
long write_matrix(const char* path, const T* data, long rows, long cols, const char* const* row_ids, const char* const* col_ids,
                  char sep, int n_threads) {
  FILE* fh = fopen(path, "wb");
  if (!fh) return -1;
  std::string head = "\"\"";
  for (long j = 0; j < cols; ++j) { head += sep; head += col_ids[j]; }
  head += '\n';
  if (fwrite(head.data(), 1, head.size(), fh) != head.size()) { fclose(fh); return -2; }
  if (n_threads < 1) n_threads = 1;
  const long block = 512;                                   // rows per work item
  const long n_blocks = (rows + block - 1) / block;
  for (long b0 = 0; b0 < n_blocks; b0 += n_threads) {       // one wave of blocks at a time keeps memory bounded
    const long nb = b0 + n_threads <= n_blocks ? n_threads : n_blocks - b0;
    std::vector<std::string> out(nb);
    std::vector<std::thread> pool;
    for (long t = 0; t < nb; ++t) {
      pool.emplace_back([&, t] {
        const long r0 = (b0 + t) * block, r1 = r0 + block < rows ? r0 + block : rows;
        std::string& s = out[t];
        s.reserve((size_t)(r1 - r0) * (size_t)(cols * 12 + 32));
        char tmp[40];
        for (long r = r0; r < r1; ++r) {
          s += row_ids[r];
          const T* row = data + r * cols;
          for (long j = 0; j < cols; ++j) {
            s += sep;
            auto res = std::to_chars(tmp, tmp + sizeof(tmp), row[j]);
            s.append(tmp, res.ptr - tmp);
          }
          s += '\n';
        }
      });
    }
    for (auto& th : pool) th.join();
    for (long t = 0; t < nb; ++t)
      if (fwrite(out[t].data(), 1, out[t].size(), fh) != out[t].size()) { fclose(fh); return -2; }
  }
  return fclose(fh) == 0 ? 0 : -2;
}

}  // namespace

extern "C" {

// data: row-major [rows][cols]; row_ids / col_ids: NUL-terminated strings written verbatim (quote them yourself if needed).
// Returns 0, -1 (cannot open) or -2 (write error).  NaN / inf are written as nan / inf / -inf.
long cml_host_csv_write_f64(const char* path, const double* data, long rows, long cols, const char* const* row_ids,
                            const char* const* col_ids, char sep, int n_threads) {
  return write_matrix<double>(path, data, rows, cols, row_ids, col_ids, sep, n_threads);
}
long cml_host_csv_write_i32(const char* path, const int32_t* data, long rows, long cols, const char* const* row_ids,
                            const char* const* col_ids, char sep, int n_threads) {
  return write_matrix<int32_t>(path, data, rows, cols, row_ids, col_ids, sep, n_threads);
}

}  // extern "C"

// Per-sample count-file concatenation: the native core of io.concat_counts.
//
// Replaces catExpnData() (reference: JSmith_code/Cat_Expn_Data.r:7-49), which read every per-sample HTSeq
// "gene<TAB>count" file with read.delim and cbind-ed one column at a time (GDC_Data_Download.Rmd:306-349; 60,488 genes x
// 474 samples).  Here every file is mmap-ed and parsed in place by a pool of threads straight into its column of the
// genes x samples int32 matrix; the gene order of the first file is the contract and is verified for every other file.
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

struct Mapped {
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
    return true;
  }
  ~Mapped() {
    if (p && n) munmap(const_cast<char*>(p), n);
    if (fd >= 0) close(fd);
  }
};

// Parses "id<sep>value" lines.  ids (if non-null) receives the id strings; ref_ids (if non-null) is checked against.
// Returns the number of rows, or -(line number) on a mismatch / parse error.
long parse_file(const Mapped& f, char sep, int skip_lines, int id_col, int val_col, std::vector<std::string>* ids,
                const std::vector<std::string>* ref_ids, int32_t* out, long stride, long max_rows) {
  const char* p = f.p;
  const char* end = f.p + f.n;
  long row = 0, line = 0;
  while (p < end) {
    const char* eol = static_cast<const char*>(memchr(p, '\n', end - p));
    if (!eol) eol = end;
    ++line;
    if (line > skip_lines && eol > p) {
      const char* q = p;
      int col = 0;
      const char* id_b = nullptr; const char* id_e = nullptr;
      long value = 0;
      bool have_val = false;
      while (q <= eol) {
        const char* fe = static_cast<const char*>(memchr(q, sep, eol - q));
        if (!fe) fe = eol;
        const char* fb = q;
        const char* fend = fe;
        if (fend > fb && fend[-1] == '\r') --fend;
        if (col == id_col) { id_b = fb; id_e = fend; }
        if (col == val_col) {
          bool neg = false;
          const char* c = fb;
          if (c < fend && *c == '-') { neg = true; ++c; }
          long v = 0;
          bool any = false;
          for (; c < fend && *c >= '0' && *c <= '9'; ++c) { v = v * 10 + (*c - '0'); any = true; }
          if (c < fend && *c == '.') {            // tolerate "12.0"
            ++c;
            while (c < fend && *c >= '0' && *c <= '9') ++c;
          }
          if (!any || c != fend) return -line;
          value = neg ? -v : v;
          have_val = true;
        }
        ++col;
        if (fe >= eol) break;
        q = fe + 1;
      }
      if (!have_val || !id_b) return -line;
      if (row >= max_rows && max_rows > 0) return -line;
      if (ids) ids->emplace_back(id_b, id_e);
      if (ref_ids) {
        const std::string& r = (*ref_ids)[row];
        if ((size_t)(id_e - id_b) != r.size() || memcmp(id_b, r.data(), r.size()) != 0) return -line;
      }
      if (out) out[row * stride] = (int32_t)value;
      ++row;
    }
    p = eol + 1;
  }
  return row;
}

std::vector<std::string> g_ids;

}  // namespace

extern "C" {

// Pass 1: read the first file to learn the gene ids.  Returns the row count (or a negative line number on error).
long cml_host_counts_scan(const char* path, char sep, int skip_lines, int id_col, int val_col) {
  Mapped f;
  if (!f.open(path)) return -1000000000L;
  g_ids.clear();
  return parse_file(f, sep, skip_lines, id_col, val_col, &g_ids, nullptr, nullptr, 0, 0);
}

// Copies gene id i (from the last scan) into buf; returns its length.
int cml_host_counts_id(long i, char* buf, int cap) {
  if (i < 0 || (size_t)i >= g_ids.size()) return -1;
  const std::string& s = g_ids[i];
  const int n = (int)s.size() < cap - 1 ? (int)s.size() : cap - 1;
  memcpy(buf, s.data(), n);
  buf[n] = 0;
  return (int)s.size();
}

// Pass 2: parse n_files files into out[genes][n_files] (int32, gene-major).  status[f] = 0 ok, otherwise the negative
// line number of the first problem (-1e9: cannot open).  Gene ids are checked against the scanned file.
int cml_host_counts_concat(const char* const* paths, int n_files, char sep, int skip_lines, int id_col, int val_col,
                           int32_t* out, long n_genes, long* status, int n_threads) {
  if (n_threads < 1) n_threads = 1;
  std::atomic<int> next{0};
  std::atomic<int> bad{0};
  auto work = [&]() {
    for (;;) {
      const int f = next.fetch_add(1);
      if (f >= n_files) break;
      Mapped m;
      if (!m.open(paths[f])) { status[f] = -1000000000L; bad++; continue; }
      const long r = parse_file(m, sep, skip_lines, id_col, val_col, nullptr, &g_ids, out + f, n_files, n_genes);
      status[f] = (r == n_genes) ? 0 : (r < 0 ? r : -(r + 1));
      if (status[f] != 0) bad++;
    }
  };
  std::vector<std::thread> pool;
  for (int t = 0; t < n_threads; ++t) pool.emplace_back(work);
  for (auto& t : pool) t.join();
  return bad.load();
}

}  // extern "C"

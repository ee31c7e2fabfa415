// Kaldi ark/scp streaming I/O (include/xvkaldi.h): the data formats either side of the embedding path, host only.
// Behaviour follows dataset/kaldi_io.py of the reference (line references in the header); the implementation is a
// buffered byte stream (files, gzip, commands) + table-driven decoders writing straight into the caller's (pinned)
// buffers, and a thread pool for scp random access.
#include <errno.h>
#include <fcntl.h>
#include <spawn.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <sys/types.h>
#include <sys/uio.h>
#include <sys/wait.h>
#include <unistd.h>
#include <zlib.h>

#include <algorithm>
#include <atomic>
#include <map>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/xvkaldi.h"

extern char** environ;

namespace {

constexpr size_t kBufBytes = 1 << 20;
constexpr size_t kDirectBytes = 16 << 10;   // bodies at least this long bypass the bounce buffer
constexpr size_t kLookahead = 4 << 10;

void set_err(std::string* e, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  *e = buf;
}

bool ends_with(const std::string& s, const char* suf) {
  const size_t n = strlen(suf);
  return s.size() >= n && s.compare(s.size() - n, n, suf) == 0;
}

std::string trim(const std::string& s) {
  size_t a = 0, b = s.size();
  while (a < b && isspace(static_cast<unsigned char>(s[a]))) ++a;
  while (b > a && isspace(static_cast<unsigned char>(s[b - 1]))) --b;
  return s.substr(a, b - a);
}

// open_or_fd (dataset/kaldi_io.py:633-640): strip "ark|scp[,opt...]:" and split a trailing ":<digits>" byte offset.
// Returns the option list (e.g. "ark,scp") in *prefix.
void parse_specifier(const std::string& spec, std::string* prefix, std::string* file, int64_t* offset) {
  *file = spec;
  prefix->clear();
  *offset = -1;
  size_t colon = spec.find(':');
  if (colon != std::string::npos && (spec.compare(0, 3, "ark") == 0 || spec.compare(0, 3, "scp") == 0)) {
    bool ok = true;
    for (size_t i = 3; i < colon; ++i) {
      const char ch = spec[i];
      if (!(ch == ',' || (ch >= 'a' && ch <= 'z'))) ok = false;
    }
    if (ok) {
      *prefix = spec.substr(0, colon);
      *file = spec.substr(colon + 1);
    }
  }
  const size_t last = file->rfind(':');
  if (last != std::string::npos && last + 1 < file->size()) {
    bool digits = true;
    for (size_t i = last + 1; i < file->size(); ++i)
      if (!isdigit(static_cast<unsigned char>((*file)[i]))) digits = false;
    if (digits) {
      *offset = strtoll(file->c_str() + last + 1, nullptr, 10);
      file->resize(last);
    }
  }
}

// posix_spawn, not fork(): the callers live in processes with CUDA contexts and busy threads (the multi-GPU extractor opens
// "ark:cmd |" while its workers create their contexts).  fork() runs every pthread_atfork handler of the process -- the CUDA
// driver's among them, which faulted when another thread was inside context creation -- and copies the page tables of a
// process that maps hundreds of GB; glibc's posix_spawn is clone(CLONE_VM | CLONE_VFORK) + exec and runs no handlers.
pid_t spawn_shell(const std::string& cmd, bool child_writes, int* fd_out, std::string* err) {
  int p[2];
  if (pipe2(p, O_CLOEXEC) != 0) { set_err(err, "pipe(): %s", strerror(errno)); return -1; }
  posix_spawn_file_actions_t fa;
  posix_spawn_file_actions_init(&fa);
  // dup2 clears FD_CLOEXEC on the target; the pipe's own descriptors close at exec
  if (child_writes) posix_spawn_file_actions_adddup2(&fa, p[1], 1);
  else posix_spawn_file_actions_adddup2(&fa, p[0], 0);
  char* const argv[] = {const_cast<char*>("sh"), const_cast<char*>("-c"), const_cast<char*>(cmd.c_str()), nullptr};
  pid_t pid = -1;
  const int rc = posix_spawn(&pid, "/bin/sh", &fa, nullptr, argv, environ);
  posix_spawn_file_actions_destroy(&fa);
  if (rc != 0) { set_err(err, "posix_spawn(): %s", strerror(rc)); close(p[0]); close(p[1]); return -1; }
  if (child_writes) { close(p[1]); *fd_out = p[0]; } else { close(p[0]); *fd_out = p[1]; }
  return pid;
}

int reap(pid_t pid, const std::string& cmd, std::string* err) {
  int status = 0;
  while (waitpid(pid, &status, 0) < 0) {
    if (errno != EINTR) { set_err(err, "waitpid(): %s", strerror(errno)); return XVK_EIO; }
  }
  if (WIFEXITED(status) && WEXITSTATUS(status) > 0) {
    set_err(err, "cmd %s returned %d !", cmd.c_str(), WEXITSTATUS(status));      // SubprocessFailed, kaldi_io.py:668-669
    return XVK_EIO;
  }
  return XVK_OK;
}

// ---------------------------------------------------------------------------------------------- byte source
struct Source {
  int fd = -1;
  gzFile gz = nullptr;
  pid_t child = -1;
  std::string cmd;
  bool seekable = false;
  std::string path;            // regular files: the name, for the per-thread descriptors of the indexed batch reader
  size_t fill_limit = 0;       // > 0: refill at most this many bytes (header scans that seek over the bodies)
  std::vector<uint8_t> buf;
  size_t pos = 0, end = 0;
  bool eof = false;
  std::string err;
  // memory-mapped mode (regular uncompressed files, batch reader): every "read" is a memcpy out of the page cache --
  // no system call per object, and the worker threads copy bodies out of the same mapping in parallel
  const uint8_t* map = nullptr;
  size_t map_size = 0, map_pos = 0;
  bool map_owner = false;

  // logical stream position (bytes consumed so far) of a seekable source
  int64_t tell() const {
    if (map) return static_cast<int64_t>(map_pos) - static_cast<int64_t>(end - pos);
    return static_cast<int64_t>(lseek(fd, 0, SEEK_CUR)) - static_cast<int64_t>(end - pos);
  }

  bool enable_map() {
    if (map) return true;
    if (!seekable || gz || fd < 0) return false;
    struct stat st;
    if (fstat(fd, &st) != 0 || st.st_size <= 0) return false;
    const int64_t at = tell();
    static const bool populate = getenv("XVK_MAP_POPULATE") != nullptr;     // A/B switch: pre-fault the whole mapping
    void* m = mmap(nullptr, static_cast<size_t>(st.st_size), PROT_READ, MAP_SHARED | (populate ? MAP_POPULATE : 0), fd, 0);
    if (m == MAP_FAILED) return false;
    madvise(m, static_cast<size_t>(st.st_size), MADV_SEQUENTIAL);
    map = static_cast<const uint8_t*>(m);
    map_size = static_cast<size_t>(st.st_size);
    map_pos = static_cast<size_t>(at);
    map_owner = true;
    pos = end = 0;
    return true;
  }

  int open(const std::string& rspecifier, size_t buf_bytes = kBufBytes) {
    std::string prefix, file;
    int64_t offset;
    parse_specifier(rspecifier, &prefix, &file, &offset);
    file = trim(file);
    if (file.empty()) { set_err(&err, "empty rspecifier"); return XVK_EOPEN; }
    buf.resize(buf_bytes);
    if (file.back() == '|') {
      cmd = trim(file.substr(0, file.size() - 1));
      child = spawn_shell(cmd, true, &fd, &err);
      if (child < 0) return XVK_EOPEN;
    } else if (ends_with(file, ".gz")) {
      gz = gzopen(file.c_str(), "rb");
      if (!gz) { set_err(&err, "cannot open %s: %s", file.c_str(), strerror(errno)); return XVK_EOPEN; }
      gzbuffer(gz, 1 << 18);
      if (offset >= 0 && gzseek(gz, offset, SEEK_SET) < 0) { set_err(&err, "cannot seek %s to %lld", file.c_str(), (long long)offset); return XVK_EOPEN; }
    } else {
      fd = ::open(file.c_str(), O_RDONLY | O_CLOEXEC);
      if (fd < 0) { set_err(&err, "cannot open %s: %s", file.c_str(), strerror(errno)); return XVK_EOPEN; }
      struct stat st;
      seekable = fstat(fd, &st) == 0 && S_ISREG(st.st_mode);
      if (seekable) path = file;
      if (offset >= 0 && lseek(fd, offset, SEEK_SET) < 0) { set_err(&err, "cannot seek %s to %lld", file.c_str(), (long long)offset); return XVK_EOPEN; }
#ifdef POSIX_FADV_SEQUENTIAL
      if (seekable && offset < 0) posix_fadvise(fd, 0, 0, POSIX_FADV_SEQUENTIAL);
#endif
    }
    return XVK_OK;
  }

  // raw read of up to n bytes; 0 at end of stream, -1 on error
  ssize_t raw_read(void* dst, size_t n) {
    if (map) {
      const size_t left = map_size - map_pos;
      const size_t take = n < left ? n : left;
      memcpy(dst, map + map_pos, take);
      map_pos += take;
      return static_cast<ssize_t>(take);
    }
    if (gz) {
      const int got = gzread(gz, dst, static_cast<unsigned>(n > (1u << 30) ? (1u << 30) : n));
      if (got < 0) { set_err(&err, "gzread failed"); return -1; }
      return got;
    }
    for (;;) {
      const ssize_t got = ::read(fd, dst, n);
      if (got < 0 && errno == EINTR) continue;
      if (got < 0) set_err(&err, "read(): %s", strerror(errno));
      return got;
    }
  }

  bool fill() {
    if (eof) return false;
    pos = 0;
    end = 0;
    const ssize_t got = raw_read(buf.data(), fill_limit ? std::min(fill_limit, buf.size()) : buf.size());
    if (got <= 0) { eof = true; return false; }
    end = static_cast<size_t>(got);
    return true;
  }

  int getc() {
    if (pos == end && !fill()) return -1;
    return buf[pos++];
  }

  // exactly n bytes, or false (short stream / error)
  bool read(void* dstv, size_t n) {
    uint8_t* dst = static_cast<uint8_t*>(dstv);
    const size_t have = end - pos;
    const size_t take = have < n ? have : n;
    memcpy(dst, buf.data() + pos, take);
    pos += take;
    dst += take;
    n -= take;
    while (n > 0) {
      if (map) {
        if (map_size - map_pos < n) { eof = true; set_err(&err, "unexpected end of stream"); return false; }
        memcpy(dst, map + map_pos, n);
        map_pos += n;
        return true;
      }
      if (n >= kDirectBytes && !gz) {
        // body: scatter read -- the rest of the body lands in the destination, a short look-ahead (the next key and
        // header) in the bounce buffer; one system call and no second copy per object
        struct iovec iov[2];
        iov[0].iov_base = dst;
        iov[0].iov_len = n;
        iov[1].iov_base = buf.data();
        iov[1].iov_len = std::min(buf.size(), kLookahead);
        ssize_t got;
        do { got = ::readv(fd, iov, 2); } while (got < 0 && errno == EINTR);
        if (got < 0) set_err(&err, "readv(): %s", strerror(errno));
        if (got <= 0) { eof = true; if (err.empty()) set_err(&err, "unexpected end of stream"); return false; }
        if (static_cast<size_t>(got) >= n) {
          pos = 0;
          end = static_cast<size_t>(got) - n;
          return true;
        }
        dst += got;
        n -= static_cast<size_t>(got);
      } else if (n >= buf.size() / 2) {          // gzip: large body straight into the destination
        const ssize_t got = raw_read(dst, n);
        if (got <= 0) { eof = true; if (err.empty()) set_err(&err, "unexpected end of stream"); return false; }
        dst += got;
        n -= static_cast<size_t>(got);
      } else {
        if (!fill()) { if (err.empty()) set_err(&err, "unexpected end of stream"); return false; }
        const size_t t = end < n ? end : n;
        memcpy(dst, buf.data(), t);
        pos = t;
        dst += t;
        n -= t;
      }
    }
    return true;
  }

  bool skip(uint64_t n) {
    const size_t have = end - pos;
    if (n <= have) { pos += static_cast<size_t>(n); return true; }
    n -= have;
    pos = end = 0;
    if (map) {
      if (map_size - map_pos < n) { set_err(&err, "unexpected end of stream"); return false; }
      map_pos += static_cast<size_t>(n);
      return true;
    }
    if (seekable && !gz) {
      struct stat st;
      const off_t cur = lseek(fd, 0, SEEK_CUR);
      if (cur >= 0 && fstat(fd, &st) == 0 && static_cast<uint64_t>(cur) + n <= static_cast<uint64_t>(st.st_size)) {
        if (lseek(fd, static_cast<off_t>(n), SEEK_CUR) >= 0) return true;
      }
      set_err(&err, "unexpected end of stream");
      return false;
    }
    while (n > 0) {
      if (!fill()) { if (err.empty()) set_err(&err, "unexpected end of stream"); return false; }
      const size_t t = end < n ? end : static_cast<size_t>(n);
      pos = t;
      n -= t;
    }
    return true;
  }

  // one text line (without the newline); false at end of stream with nothing read
  bool read_line(std::string* line) {
    line->clear();
    bool any = false;
    for (;;) {
      if (pos == end && !fill()) return any;
      any = true;
      const uint8_t* p = buf.data() + pos;
      const void* nl = memchr(p, '\n', end - pos);
      if (nl) {
        const size_t len = static_cast<const uint8_t*>(nl) - p;
        line->append(reinterpret_cast<const char*>(p), len);
        pos += len + 1;
        return true;
      }
      line->append(reinterpret_cast<const char*>(p), end - pos);
      pos = end;
    }
  }

  int close_all() {
    int rc = XVK_OK;
    if (map && map_owner) munmap(const_cast<uint8_t*>(map), map_size);
    map = nullptr;
    if (gz) { gzclose(gz); gz = nullptr; }
    if (fd >= 0) { ::close(fd); fd = -1; }
    if (child > 0) { rc = reap(child, cmd, &err); child = -1; }
    return rc;
  }
};

// ---------------------------------------------------------------------------------------------- object decoding
struct ObjHeader {
  int32_t kind = 0, rows = 0, cols = 0;
  float cm_min = 0.f, cm_range = 0.f;
  std::vector<double> text;            // values of a text object (parsed with the header)
};

int32_t load_i32(const uint8_t* p) { int32_t v; memcpy(&v, p, 4); return v; }

// Text body after " [" (dataset/kaldi_io.py:1056-1068): rows are lines, the line that ends with "]" is the last.
int parse_text_body(Source& s, ObjHeader* h, std::string* err) {
  std::string line;
  h->rows = 0;
  h->cols = -1;
  for (;;) {
    if (!s.read_line(&line)) { set_err(err, "BadInputFormat: end of stream inside a text matrix"); return XVK_EFORMAT; }
    const char* p = line.c_str();
    int n = 0;
    bool last = false;
    for (;;) {
      while (*p && isspace(static_cast<unsigned char>(*p))) ++p;
      if (!*p) break;
      if (*p == ']') { last = true; break; }
      char* q = nullptr;
      const double v = strtod(p, &q);
      if (q == p) { set_err(err, "BadInputFormat: cannot parse '%s' in a text matrix", p); return XVK_EFORMAT; }
      h->text.push_back(v);
      ++n;
      p = q;
    }
    if (n == 0 && !last) continue;            // empty line
    if (n > 0) {
      if (h->cols < 0) h->cols = n;
      else if (h->cols != n) { set_err(err, "BadInputFormat: text matrix rows of %d and %d values", h->cols, n); return XVK_EFORMAT; }
      ++h->rows;
    }
    if (last) break;
  }
  if (h->cols < 0) h->cols = 0;
  h->kind = XVK_MAT_TEXT;
  return XVK_OK;
}

// After the key: "\0B" + token + dimensions, or " [" + text.
int parse_header(Source& s, ObjHeader* h, std::string* err) {
  uint8_t b[16];
  h->text.clear();
  if (!s.read(b, 2)) { *err = s.err; return XVK_EIO; }
  if (b[0] == 0 && b[1] == 'B') {
    if (!s.read(b, 3)) { *err = s.err; return XVK_EIO; }
    const std::string tok(reinterpret_cast<const char*>(b), 3);
    if (tok == "FM " || tok == "DM ") {
      if (!s.read(b, 10)) { *err = s.err; return XVK_EIO; }
      h->kind = tok[0] == 'F' ? XVK_MAT_F32 : XVK_MAT_F64;
      h->rows = load_i32(b + 1);
      h->cols = load_i32(b + 6);
    } else if (tok == "FV " || tok == "DV ") {
      if (!s.read(b, 5)) { *err = s.err; return XVK_EIO; }
      if (b[0] != 4) { set_err(err, "vector dimension is not preceded by the int-size marker \\4"); return XVK_EFORMAT; }
      h->kind = tok[0] == 'F' ? XVK_VEC_F32 : XVK_VEC_F64;
      h->rows = 1;
      h->cols = load_i32(b + 1);
    } else if (tok.compare(0, 2, "CM") == 0) {
      if (tok != "CM ") { set_err(err, "compressed matrix format '%s': the formats CM2, CM3 are not supported", tok.c_str()); return XVK_EFORMAT; }
      if (!s.read(b, 16)) { *err = s.err; return XVK_EIO; }
      memcpy(&h->cm_min, b, 4);
      memcpy(&h->cm_range, b + 4, 4);
      h->kind = XVK_MAT_CM;
      h->rows = load_i32(b + 8);
      h->cols = load_i32(b + 12);
    } else {
      set_err(err, "UnknownMatrixHeader: The header contained '%s'", tok.c_str());
      return XVK_EFORMAT;
    }
    if (h->rows < 0 || h->cols < 0) { set_err(err, "negative dimension %d x %d", h->rows, h->cols); return XVK_EFORMAT; }
    return XVK_OK;
  }
  if (b[0] == ' ' && b[1] == '[') return parse_text_body(s, h, err);
  set_err(err, "BadInputFormat: expected \"\\0B\" or \" [\" after the key, got bytes %02x %02x", b[0], b[1]);
  return XVK_EFORMAT;
}

uint64_t body_bytes(const ObjHeader& h) {
  const uint64_t n = static_cast<uint64_t>(h.rows) * static_cast<uint64_t>(h.cols);
  switch (h.kind) {
    case XVK_MAT_F32: case XVK_VEC_F32: return n * 4;
    case XVK_MAT_F64: case XVK_VEC_F64: return n * 8;
    case XVK_MAT_CM: return static_cast<uint64_t>(h.cols) * 8 + n;
    default: return 0;
  }
}

// CM element tables (dataset/kaldi_io.py:1083-1113), evaluated in float32 in the reference's operation order:
//   percentile = min + range * 1.52590218966964e-05 * u16
//   v <= 64: p0 + (p25-p0)/64 * v ; v <= 192: p25 + (p75-p25)/128 * (v-64) ; else p75 + (p100-p75)/63 * (v-192)
void cm_column_table(float gmin, float grange, const uint8_t* col_header, float* lut) {
  const float c = static_cast<float>(1.52590218966964e-05);
  float p[4];
  for (int i = 0; i < 4; ++i) {
    uint16_t u;
    memcpy(&u, col_header + 2 * i, 2);
    const float t = grange * c;
    const float t2 = t * static_cast<float>(u);
    p[i] = gmin + t2;
  }
  const float q0 = (p[1] - p[0]) / 64.f, q1 = (p[2] - p[1]) / 128.f, q2 = (p[3] - p[2]) / 63.f;
  for (int v = 0; v <= 64; ++v) { const float m = q0 * static_cast<float>(v); lut[v] = p[0] + m; }
  for (int v = 65; v <= 192; ++v) { const float m = q1 * static_cast<float>(v - 64); lut[v] = p[1] + m; }
  for (int v = 193; v <= 255; ++v) { const float m = q2 * static_cast<float>(v - 192); lut[v] = p[2] + m; }
}

template <class T>
int decode_body(Source& s, const ObjHeader& h, int32_t start, int32_t length, T* dst, int64_t ld, std::vector<uint8_t>* scratch,
                std::string* err) {
  const int64_t rows = h.rows, cols = h.cols;
  if (start < 0 || length < 0 || static_cast<int64_t>(start) + length > rows) {
    set_err(err, "The number of frames is not enough for length %d", length);       // kaldi_io.py:1046
    return XVK_EINVAL;
  }
  if (ld < cols) { set_err(err, "row stride %lld is smaller than the %lld columns", (long long)ld, (long long)cols); return XVK_EINVAL; }
  const bool whole = start == 0 && length == rows;
  switch (h.kind) {
    case XVK_MAT_F32: case XVK_VEC_F32: case XVK_MAT_F64: case XVK_VEC_F64: {
      const bool f32 = h.kind == XVK_MAT_F32 || h.kind == XVK_VEC_F32;
      const size_t esz = f32 ? 4 : 8;
      if (!s.skip(static_cast<uint64_t>(start) * cols * esz)) { *err = s.err; return XVK_EIO; }
      if (f32 == (sizeof(T) == 4) && ld == cols) {
        if (!s.read(dst, static_cast<size_t>(length) * cols * esz)) { *err = s.err; return XVK_EIO; }
      } else if (f32 == (sizeof(T) == 4)) {
        for (int64_t r = 0; r < length; ++r)
          if (!s.read(dst + r * ld, static_cast<size_t>(cols) * esz)) { *err = s.err; return XVK_EIO; }
      } else {
        const int64_t chunk_rows = cols > 0 ? std::max<int64_t>(1, static_cast<int64_t>(1 << 16) / cols) : 1;
        scratch->resize(static_cast<size_t>(chunk_rows) * cols * esz);
        for (int64_t r0 = 0; r0 < length; r0 += chunk_rows) {
          const int64_t nr = std::min<int64_t>(chunk_rows, length - r0);
          if (!s.read(scratch->data(), static_cast<size_t>(nr) * cols * esz)) { *err = s.err; return XVK_EIO; }
          for (int64_t r = 0; r < nr; ++r) {
            T* out = dst + (r0 + r) * ld;
            if (f32) {
              const float* in = reinterpret_cast<const float*>(scratch->data()) + r * cols;
              for (int64_t c = 0; c < cols; ++c) out[c] = static_cast<T>(in[c]);
            } else {
              const double* in = reinterpret_cast<const double*>(scratch->data()) + r * cols;
              for (int64_t c = 0; c < cols; ++c) out[c] = static_cast<T>(in[c]);
            }
          }
        }
      }
      if (!whole && !s.skip(static_cast<uint64_t>(rows - start - length) * cols * esz)) { *err = s.err; return XVK_EIO; }
      return XVK_OK;
    }
    case XVK_MAT_CM: {
      const size_t hdr = static_cast<size_t>(cols) * 8, data = static_cast<size_t>(cols) * rows;
      const size_t lut_off = (hdr + data + 15) / 16 * 16;
      scratch->resize(lut_off + static_cast<size_t>(cols) * 256 * sizeof(float));
      uint8_t* base = scratch->data();
      if (!s.read(base, hdr + data)) { *err = s.err; return XVK_EIO; }
      float* lut = reinterpret_cast<float*>(base + lut_off);
      for (int64_t c = 0; c < cols; ++c) cm_column_table(h.cm_min, h.cm_range, base + 8 * c, lut + 256 * c);
      const uint8_t* bytes = base + hdr;
      constexpr int64_t kBlock = 64;                      // rows per pass: the destination block stays in L1
      for (int64_t r0 = 0; r0 < length; r0 += kBlock) {
        const int64_t nr = std::min<int64_t>(kBlock, length - r0);
        for (int64_t c = 0; c < cols; ++c) {
          const uint8_t* src = bytes + c * rows + start + r0;
          const float* t = lut + 256 * c;
          T* out = dst + r0 * ld + c;
          for (int64_t r = 0; r < nr; ++r) out[r * ld] = static_cast<T>(t[src[r]]);
        }
      }
      return XVK_OK;
    }
    case XVK_MAT_TEXT: case XVK_VEC_TEXT: {
      for (int64_t r = 0; r < length; ++r)
        for (int64_t c = 0; c < cols; ++c) {
          const double v = h.text[static_cast<size_t>((start + r) * cols + c)];
          dst[r * ld + c] = sizeof(T) == 4 ? static_cast<T>(static_cast<float>(v)) : static_cast<T>(v);
        }
      return XVK_OK;
    }
    default:
      set_err(err, "no pending object");
      return XVK_ESTATE;
  }
}

}  // namespace

// ================================================================================================ ark reader
struct xvk_reader {
  Source src;
  int threads = 1;
  std::vector<int> worker_fds;
  bool keyless = false, first = true, pending = false;
  std::string key, err;
  ObjHeader h;
  std::vector<uint8_t> scratch;
};

namespace {

void copy_err(char* err, size_t cap, const std::string& msg) {
  if (err && cap) snprintf(err, cap, "%s", msg.c_str());
}

// read_key (dataset/kaldi_io.py:693-707): bytes up to the first ' ' (or end of stream), stripped; "" = end of stream.
int read_key(Source& s, std::string* key, std::string* err) {
  key->clear();
  for (;;) {
    if (s.pos == s.end && !s.fill()) break;
    const uint8_t* p = s.buf.data() + s.pos;
    const void* sp = memchr(p, ' ', s.end - s.pos);
    if (sp) {
      const size_t len = static_cast<const uint8_t*>(sp) - p;
      key->append(reinterpret_cast<const char*>(p), len);
      s.pos += len + 1;
      break;
    }
    key->append(reinterpret_cast<const char*>(p), s.end - s.pos);
    s.pos = s.end;
  }
  if (!s.err.empty()) { *err = s.err; return XVK_EIO; }
  *key = trim(*key);
  if (key->empty()) return 0;
  for (const char ch : *key)
    if (isspace(static_cast<unsigned char>(ch))) { set_err(err, "key '%s' contains whitespace", key->c_str()); return XVK_EFORMAT; }
  return 1;
}

int skip_pending(xvk_reader* r) {
  if (!r->pending) return XVK_OK;
  r->pending = false;
  if (!r->src.skip(body_bytes(r->h))) { r->err = r->src.err; return XVK_EIO; }
  return XVK_OK;
}

template <class Fn>
int parallel_for(int64_t n, int threads, Fn fn, std::string* err);

// Regular files: scan the headers of a batch (one short read + one seek per object), then decode the bodies with
// `threads` workers, each through its own descriptor positioned on the body.
int read_batch_indexed(xvk_reader* r, float* dst, int64_t cap_rows, int32_t dim, int32_t max_utts, int32_t* lengths,
                       char* keys, size_t keys_cap, int32_t* n_utts, int64_t* n_rows) {
  struct Item { int64_t off, row0; ObjHeader h; };
  std::vector<Item> items;
  int64_t rows = 0;
  size_t kpos = 0;
  int more = 1;
  *n_utts = 0;
  *n_rows = 0;
  Source& s = r->src;
  static const bool no_map = getenv("XVK_NO_MMAP") != nullptr;      // A/B switch
  if (!no_map) s.enable_map();
  s.fill_limit = s.map ? 256 : 4096;          // header scans: a key + header is a few dozen bytes
  for (;;) {
    if (!r->pending) {
      const int rc = xvk_next(r, nullptr, nullptr, nullptr, nullptr);
      if (rc < 0) { s.fill_limit = 0; return rc; }
      if (rc == 0) { more = 0; break; }
    }
    const ObjHeader& h = r->h;
    int bad = XVK_OK;
    if (h.kind == XVK_VEC_F32 || h.kind == XVK_VEC_F64 || h.kind == XVK_VEC_TEXT) {
      set_err(&r->err, "key %s: expected a feature matrix, found a vector", r->key.c_str());
      bad = XVK_EFORMAT;
    } else if (h.cols != dim && h.rows > 0) {
      set_err(&r->err, "key %s: feature dimension %d, expected %d", r->key.c_str(), h.cols, dim);
      bad = XVK_EFORMAT;
    } else if (static_cast<int32_t>(items.size()) == max_utts || rows + h.rows > cap_rows || kpos + r->key.size() + 1 > keys_cap) {
      if (items.empty()) {
        set_err(&r->err, "key %s (%d rows) does not fit an empty batch of %lld rows", r->key.c_str(), h.rows, (long long)cap_rows);
        bad = XVK_ECAPACITY;
      } else {
        break;
      }
    }
    if (bad != XVK_OK) { s.fill_limit = 0; return bad; }
    Item it;
    it.off = s.tell();
    it.row0 = rows;
    it.h = h;
    items.push_back(it);
    memcpy(keys + kpos, r->key.c_str(), r->key.size() + 1);
    kpos += r->key.size() + 1;
    lengths[items.size() - 1] = h.rows;
    rows += h.rows;
    r->pending = false;
    if (!s.skip(body_bytes(h))) { r->err = s.err; s.fill_limit = 0; return XVK_EIO; }
  }
  s.fill_limit = 0;
  const int threads = std::min<int64_t>(r->threads, std::max<int64_t>(1, static_cast<int64_t>(items.size())));
  if (s.map) {
    // bodies are copied (FM) or decoded (DM / CM) straight out of the mapping, one object per task
    const int mrc = parallel_for(static_cast<int64_t>(items.size()), threads, [&](int64_t i, int, std::vector<uint8_t>* scratch, std::string* err) {
      const Item& it = items[i];
      Source w;
      w.map = s.map;
      w.map_size = s.map_size;
      w.map_pos = static_cast<size_t>(it.off);
      w.seekable = true;
      return decode_body(w, it.h, 0, it.h.rows, dst + it.row0 * dim, static_cast<int64_t>(dim), scratch, err);
    }, &r->err);
    if (mrc != XVK_OK) return mrc;
    *n_utts = static_cast<int32_t>(items.size());
    *n_rows = rows;
    return more;
  }
  while (static_cast<int>(r->worker_fds.size()) < threads) {
    const int fd = ::open(s.path.c_str(), O_RDONLY | O_CLOEXEC);
    if (fd < 0) { set_err(&r->err, "cannot open %s: %s", s.path.c_str(), strerror(errno)); return XVK_EOPEN; }
    r->worker_fds.push_back(fd);
  }
  const int rc = parallel_for(static_cast<int64_t>(items.size()), threads, [&](int64_t i, int tid, std::vector<uint8_t>* scratch, std::string* err) {
    const Item& it = items[i];
    Source w;
    w.fd = r->worker_fds[tid];
    w.seekable = true;
    w.buf.resize(1 << 12);
    if (lseek(w.fd, it.off, SEEK_SET) < 0) { set_err(err, "lseek(): %s", strerror(errno)); w.fd = -1; return static_cast<int>(XVK_EIO); }
    const int drc = decode_body(w, it.h, 0, it.h.rows, dst + it.row0 * dim, static_cast<int64_t>(dim), scratch, err);
    w.fd = -1;                                  // the descriptor belongs to the reader
    return drc;
  }, &r->err);
  if (rc != XVK_OK) return rc;
  *n_utts = static_cast<int32_t>(items.size());
  *n_rows = rows;
  return more;
}

template <class T>
int reader_decode(xvk_reader* r, int32_t start, int32_t length, T* dst, int64_t ld) {
  if (!r) return XVK_EINVAL;
  if (!r->pending) { r->err = "no pending object: call xvk_next first"; return XVK_ESTATE; }
  if (!dst && static_cast<int64_t>(length) * r->h.cols > 0) { r->err = "null destination"; return XVK_EINVAL; }
  const int rc = decode_body(r->src, r->h, start, length, dst, ld, &r->scratch, &r->err);
  if (rc == XVK_EINVAL) return rc;        // nothing consumed: the object stays pending
  r->pending = false;
  return rc;
}

}  // namespace

extern "C" {

xvk_reader* xvk_reader_open(const char* rspecifier, int keyless, char* err, size_t errcap) {
  if (!rspecifier) { copy_err(err, errcap, "null rspecifier"); return nullptr; }
  xvk_reader* r = new xvk_reader();
  r->keyless = keyless != 0;
  if (r->src.open(rspecifier) != XVK_OK) {
    copy_err(err, errcap, r->src.err);
    r->src.close_all();
    delete r;
    return nullptr;
  }
  return r;
}

int xvk_next(xvk_reader* r, const char** key, int32_t* kind, int32_t* rows, int32_t* cols) {
  if (!r) return XVK_EINVAL;
  int rc = skip_pending(r);
  if (rc != XVK_OK) return rc;
  if (r->keyless) {
    if (!r->first) return 0;
    r->key.clear();
  } else {
    rc = read_key(r->src, &r->key, &r->err);
    if (rc <= 0) return rc;
  }
  r->first = false;
  rc = parse_header(r->src, &r->h, &r->err);
  if (rc != XVK_OK) return rc;
  r->pending = true;
  if (key) *key = r->key.c_str();
  if (kind) *kind = r->h.kind;
  if (rows) *rows = r->h.rows;
  if (cols) *cols = r->h.cols;
  return 1;
}

int xvk_read_f32(xvk_reader* r, float* dst, int64_t ld) { return reader_decode(r, 0, r ? r->h.rows : 0, dst, ld); }
int xvk_read_f64(xvk_reader* r, double* dst, int64_t ld) { return reader_decode(r, 0, r ? r->h.rows : 0, dst, ld); }
int xvk_read_rows_f32(xvk_reader* r, int32_t start, int32_t length, float* dst, int64_t ld) {
  return reader_decode(r, start, length, dst, ld);
}

int xvk_reader_set_threads(xvk_reader* r, int32_t threads) {
  if (!r || threads < 1) return XVK_EINVAL;
  r->threads = threads;
  return XVK_OK;
}

int xvk_read_batch_f32(xvk_reader* r, float* dst, int64_t cap_rows, int32_t dim, int32_t max_utts, int32_t* lengths,
                       char* keys, size_t keys_cap, int32_t* n_utts, int64_t* n_rows) {
  if (!r || !dst || !lengths || !keys || !n_utts || !n_rows || dim <= 0 || max_utts <= 0) return XVK_EINVAL;
  if (r->threads > 1 && r->src.seekable && !r->src.gz && !r->src.path.empty())
    return read_batch_indexed(r, dst, cap_rows, dim, max_utts, lengths, keys, keys_cap, n_utts, n_rows);
  int32_t n = 0;
  int64_t rows = 0;
  size_t kpos = 0;
  *n_utts = 0;
  *n_rows = 0;
  for (;;) {
    if (!r->pending) {
      const int rc = xvk_next(r, nullptr, nullptr, nullptr, nullptr);
      if (rc < 0) return rc;
      if (rc == 0) { *n_utts = n; *n_rows = rows; return 0; }
    }
    const ObjHeader& h = r->h;
    if (h.kind == XVK_VEC_F32 || h.kind == XVK_VEC_F64 || h.kind == XVK_VEC_TEXT) {
      set_err(&r->err, "key %s: expected a feature matrix, found a vector", r->key.c_str());
      return XVK_EFORMAT;
    }
    if (h.cols != dim && h.rows > 0) {
      set_err(&r->err, "key %s: feature dimension %d, expected %d", r->key.c_str(), h.cols, dim);
      return XVK_EFORMAT;
    }
    if (n == max_utts || rows + h.rows > cap_rows || kpos + r->key.size() + 1 > keys_cap) {
      if (n == 0) {
        set_err(&r->err, "key %s (%d rows) does not fit an empty batch of %lld rows", r->key.c_str(), h.rows, (long long)cap_rows);
        return XVK_ECAPACITY;
      }
      *n_utts = n;
      *n_rows = rows;
      return 1;
    }
    const int rc = reader_decode(r, 0, h.rows, dst + rows * dim, dim);
    if (rc != XVK_OK) return rc;
    memcpy(keys + kpos, r->key.c_str(), r->key.size() + 1);
    kpos += r->key.size() + 1;
    lengths[n++] = h.rows;
    rows += h.rows;
  }
}

const char* xvk_reader_error(const xvk_reader* r) { return r ? r->err.c_str() : "null reader"; }

int xvk_reader_close(xvk_reader* r) {
  if (!r) return XVK_OK;
  for (const int fd : r->worker_fds) ::close(fd);
  const int rc = r->src.close_all();
  delete r;
  return rc;
}

}  // extern "C"

// ================================================================================================ scp
struct ScpEntry {
  std::string key, rxfile;
};

struct xvk_scp {
  std::vector<ScpEntry> entries;
  std::string err;
  std::vector<uint8_t> scratch;
  // "file:offset" targets in regular uncompressed files are served out of one read-only mapping per file, created on first
  // use and kept until the table is closed: a random-access read is then a header parse and a memcpy -- no open / lseek /
  // read / close per entry (a training batch is 64 such reads per millisecond).  p == nullptr: the file cannot be mapped.
  struct FileMap { const uint8_t* p; size_t size; };
  std::mutex map_mu;
  std::map<std::string, FileMap> maps;
  ~xvk_scp() {
    for (auto& kv : maps)
      if (kv.second.p) munmap(const_cast<uint8_t*>(kv.second.p), kv.second.size);
  }
};

namespace {

// One scp target: open (with its byte offset), parse the header; optionally decode.
// start < 0: the whole matrix; else rows [start, start + length) of it (FeatureReader.read_segment, dataset/kaldi_io.py:1034-1053)
// The mapping of an scp target's file (see xvk_scp::maps); false: not a plain "file:offset" target or not mappable.
bool scp_mapping(xvk_scp* owner, const std::string& rxfile, xvk_scp::FileMap* out, int64_t* offset) {
  static const bool no_map = getenv("XVK_NO_MMAP") != nullptr;      // A/B switch
  if (!owner || no_map) return false;
  std::string prefix, file;
  parse_specifier(rxfile, &prefix, &file, offset);
  file = trim(file);
  if (*offset < 0 || file.empty() || file.back() == '|' || ends_with(file, ".gz")) return false;
  std::lock_guard<std::mutex> lock(owner->map_mu);
  auto it = owner->maps.find(file);
  if (it == owner->maps.end()) {
    xvk_scp::FileMap m = {nullptr, 0};
    const int fd = ::open(file.c_str(), O_RDONLY | O_CLOEXEC);
    if (fd >= 0) {
      struct stat st;
      if (fstat(fd, &st) == 0 && S_ISREG(st.st_mode) && st.st_size > 0) {
        void* p = mmap(nullptr, static_cast<size_t>(st.st_size), PROT_READ, MAP_SHARED, fd, 0);
        if (p != MAP_FAILED) {
          madvise(p, static_cast<size_t>(st.st_size), MADV_RANDOM);
          m.p = static_cast<const uint8_t*>(p);
          m.size = static_cast<size_t>(st.st_size);
        }
      }
      ::close(fd);
    }
    it = owner->maps.emplace(file, m).first;
  }
  if (!it->second.p || static_cast<size_t>(*offset) >= it->second.size) return false;
  *out = it->second;
  return true;
}

int scp_fetch(xvk_scp* owner, const ScpEntry& e, ObjHeader* h, float* dst, int64_t ld, int64_t cap_rows, std::vector<uint8_t>* scratch,
              std::string* err, int32_t start = -1, int32_t length = 0) {
  Source s;
  xvk_scp::FileMap fm;
  int64_t offset = -1;
  int rc = XVK_OK;
  if (scp_mapping(owner, e.rxfile, &fm, &offset)) {
    s.map = fm.p;
    s.map_size = fm.size;
    s.map_pos = static_cast<size_t>(offset);
    s.seekable = true;                           // map_owner stays false: the table owns the mapping
    s.buf.resize(256);                           // look-ahead of the header parse (a token and two int32 fields)
  } else {
    rc = s.open(e.rxfile, 1 << 16);              // random access: small read-ahead
    if (rc != XVK_OK) { *err = s.err; s.close_all(); return rc; }
  }
  rc = parse_header(s, h, err);
  if (rc == XVK_OK && dst != nullptr) {
    const int32_t first = start < 0 ? 0 : start;
    const int32_t count = start < 0 ? h->rows : length;
    if (count > cap_rows) {
      set_err(err, "key %s: %d rows do not fit the %lld rows left in the destination", e.key.c_str(), count, (long long)cap_rows);
      rc = XVK_ECAPACITY;
    } else {
      rc = decode_body(s, *h, first, count, dst, ld, scratch, err);
      if (rc != XVK_OK && err->find("key ") != 0) *err = "key " + e.key + ": " + *err;
    }
  }
  const int rc2 = s.close_all();
  if (rc == XVK_OK && rc2 != XVK_OK) { *err = s.err; rc = rc2; }
  return rc;
}

// Parked worker threads shared by every reader of the process: a batch of the ark reader and the segments of a training batch
// are ~1 ms jobs, and starting 8-32 threads for each of them cost a quarter of that.  One job at a time (a caller that finds
// the pool busy starts its own threads as before); the caller works as thread 0.  A forked child finds the parent's pid
// recorded and starts over with no workers (the parent's threads do not exist there).
class WorkerPool {
  // Everything the workers touch lives in one heap object: a forked child cannot reuse the parent's (its condition variables
  // were copied with waiters that do not exist in the child), so it abandons that object and starts a fresh one.
  struct State {
    std::mutex mu;
    std::condition_variable cv_start, cv_done;
    std::vector<std::thread> workers;
    const std::function<void(int)>* job = nullptr;
    int job_n = 0, pending = 0;
    uint64_t gen = 0;
    bool stop = false;
  };

 public:
  bool run(int n, const std::function<void(int)>& fn) {
    std::unique_lock<std::mutex> use(use_mu_, std::try_to_lock);
    if (!use.owns_lock()) return false;
    if (owner_pid_ != getpid()) {                         // first use, or a forked child (the old state is leaked on purpose)
      st_ = new State();
      owner_pid_ = getpid();
    }
    State* st = st_;
    while (static_cast<int>(st->workers.size()) < n - 1) {
      const int w = static_cast<int>(st->workers.size()) + 1;
      st->workers.emplace_back([st, w] { loop(st, w); });
    }
    {
      std::lock_guard<std::mutex> lock(st->mu);
      st->job = &fn;
      st->job_n = n;
      st->pending = n - 1;
      ++st->gen;
    }
    st->cv_start.notify_all();
    fn(0);
    std::unique_lock<std::mutex> lock(st->mu);
    st->cv_done.wait(lock, [st] { return st->pending == 0; });
    st->job = nullptr;
    return true;
  }
  ~WorkerPool() {
    if (!st_ || owner_pid_ != getpid()) return;           // never used here, or a forked child: nothing to join
    {
      std::lock_guard<std::mutex> lock(st_->mu);
      st_->stop = true;
    }
    st_->cv_start.notify_all();
    for (auto& t : st_->workers) t.join();
    delete st_;
  }

 private:
  static void loop(State* st, int w) {
    uint64_t seen = 0;
    for (;;) {
      const std::function<void(int)>* job = nullptr;
      {
        std::unique_lock<std::mutex> lock(st->mu);
        st->cv_start.wait(lock, [&] { return st->stop || st->gen != seen; });
        if (st->stop) return;
        seen = st->gen;
        if (w < st->job_n) job = st->job;
      }
      if (!job) continue;
      (*job)(w);
      std::lock_guard<std::mutex> lock(st->mu);
      if (--st->pending == 0) st->cv_done.notify_one();
    }
  }
  std::mutex use_mu_;
  State* st_ = nullptr;
  pid_t owner_pid_ = -1;
};
WorkerPool g_pool;

template <class Fn>
int parallel_for(int64_t n, int threads, Fn fn, std::string* err) {
  if (threads < 1) threads = 1;
  if (threads > n) threads = static_cast<int>(n > 0 ? n : 1);
  std::atomic<int64_t> next(0);
  std::atomic<int> status(XVK_OK);
  std::vector<std::string> errs(threads);
  const std::function<void(int)> work = [&](int tid) {
    static thread_local std::vector<uint8_t> scratch;       // decode scratch of this thread, kept across jobs
    for (;;) {
      const int64_t i = next.fetch_add(1);
      if (i >= n || status.load() != XVK_OK) return;
      const int rc = fn(i, tid, &scratch, &errs[tid]);
      if (rc != XVK_OK) { int expect = XVK_OK; status.compare_exchange_strong(expect, rc); return; }
    }
  };
  static const bool no_pool = getenv("XVK_NO_POOL") != nullptr;      // A/B switch: a fresh set of threads per job
  if (threads == 1) {
    work(0);
  } else if (no_pool || !g_pool.run(threads, work)) {
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t) pool.emplace_back(work, t);
    for (auto& th : pool) th.join();
  }
  if (status.load() != XVK_OK)
    for (const auto& m : errs)
      if (!m.empty()) { *err = m; break; }
  return status.load();
}

}  // namespace

extern "C" {

xvk_scp* xvk_scp_open(const char* path, char* err, size_t errcap) {
  if (!path) { copy_err(err, errcap, "null scp path"); return nullptr; }
  Source s;
  if (s.open(path) != XVK_OK) { copy_err(err, errcap, s.err); s.close_all(); return nullptr; }
  xvk_scp* scp = new xvk_scp();
  std::string line;
  int64_t lineno = 0;
  while (s.read_line(&line)) {
    ++lineno;
    const std::string t = trim(line);
    if (t.empty()) continue;
    const size_t sp = t.find_first_of(" \t");
    if (sp == std::string::npos) {
      std::string m;
      set_err(&m, "%s line %lld: expected \"key rxfile\"", path, (long long)lineno);
      copy_err(err, errcap, m);
      s.close_all();
      delete scp;
      return nullptr;
    }
    ScpEntry e;
    e.key = t.substr(0, sp);
    e.rxfile = trim(t.substr(sp + 1));
    scp->entries.push_back(e);
  }
  const int rc = s.close_all();
  if (rc != XVK_OK) { copy_err(err, errcap, s.err); delete scp; return nullptr; }
  return scp;
}

int64_t xvk_scp_count(const xvk_scp* s) { return s ? static_cast<int64_t>(s->entries.size()) : 0; }
const char* xvk_scp_key(const xvk_scp* s, int64_t i) {
  return (s && i >= 0 && i < static_cast<int64_t>(s->entries.size())) ? s->entries[i].key.c_str() : nullptr;
}
const char* xvk_scp_rxfile(const xvk_scp* s, int64_t i) {
  return (s && i >= 0 && i < static_cast<int64_t>(s->entries.size())) ? s->entries[i].rxfile.c_str() : nullptr;
}

int xvk_scp_dims(xvk_scp* s, int64_t i, int32_t* kind, int32_t* rows, int32_t* cols) {
  if (!s || i < 0 || i >= static_cast<int64_t>(s->entries.size())) return XVK_EINVAL;
  ObjHeader h;
  const int rc = scp_fetch(s, s->entries[i], &h, nullptr, 0, 0, &s->scratch, &s->err);
  if (rc != XVK_OK) return rc;
  if (kind) *kind = h.kind;
  if (rows) *rows = h.rows;
  if (cols) *cols = h.cols;
  return XVK_OK;
}

int xvk_scp_read_f32(xvk_scp* s, int64_t i, float* dst, int64_t ld) {
  if (!s || !dst || i < 0 || i >= static_cast<int64_t>(s->entries.size())) return XVK_EINVAL;
  ObjHeader h;
  return scp_fetch(s, s->entries[i], &h, dst, ld, INT64_MAX, &s->scratch, &s->err);
}

int xvk_scp_read_many_f32(xvk_scp* s, const int64_t* idx, int64_t n, float* dst, int64_t ld, const int64_t* row_off,
                          int64_t cap_rows, int32_t* rows_out, int32_t threads) {
  if (!s || !idx || !dst || !row_off || n < 0) return XVK_EINVAL;
  const int64_t count = static_cast<int64_t>(s->entries.size());
  for (int64_t k = 0; k < n; ++k)
    if (idx[k] < 0 || idx[k] >= count || row_off[k] < 0 || row_off[k] > cap_rows) { s->err = "index or row offset out of range"; return XVK_EINVAL; }
  return parallel_for(n, threads, [&](int64_t k, int, std::vector<uint8_t>* scratch, std::string* err) {
    ObjHeader h;
    const int rc = scp_fetch(s, s->entries[idx[k]], &h, dst + row_off[k] * ld, ld, cap_rows - row_off[k], scratch, err);
    if (rc == XVK_OK && rows_out) rows_out[k] = h.rows;
    return rc;
  }, &s->err);
}

int xvk_scp_read_segments_f32(xvk_scp* s, const int64_t* idx, const int32_t* start, const int32_t* length, int64_t n, float* dst,
                              int64_t ld, const int64_t* row_off, int64_t cap_rows, int32_t threads) {
  if (!s || !idx || !start || !length || !dst || !row_off || n < 0) return XVK_EINVAL;
  const int64_t count = static_cast<int64_t>(s->entries.size());
  for (int64_t k = 0; k < n; ++k)
    if (idx[k] < 0 || idx[k] >= count || start[k] < 0 || length[k] < 0 || row_off[k] < 0 || row_off[k] > cap_rows) {
      s->err = "index, segment or row offset out of range";
      return XVK_EINVAL;
    }
  return parallel_for(n, threads, [&](int64_t k, int, std::vector<uint8_t>* scratch, std::string* err) {
    ObjHeader h;
    return scp_fetch(s, s->entries[idx[k]], &h, dst + row_off[k] * ld, ld, cap_rows - row_off[k], scratch, err, start[k], length[k]);
  }, &s->err);
}

int xvk_scp_dims_many(xvk_scp* s, const int64_t* idx, int64_t n, int32_t* rows_out, int32_t* cols_out, int32_t threads) {
  if (!s || !idx || n < 0) return XVK_EINVAL;
  const int64_t count = static_cast<int64_t>(s->entries.size());
  for (int64_t k = 0; k < n; ++k)
    if (idx[k] < 0 || idx[k] >= count) { s->err = "index out of range"; return XVK_EINVAL; }
  return parallel_for(n, threads, [&](int64_t k, int, std::vector<uint8_t>* scratch, std::string* err) {
    ObjHeader h;
    const int rc = scp_fetch(s, s->entries[idx[k]], &h, nullptr, 0, 0, scratch, err);
    if (rc == XVK_OK) {
      if (rows_out) rows_out[k] = h.rows;
      if (cols_out) cols_out[k] = h.cols;
    }
    return rc;
  }, &s->err);
}

const char* xvk_scp_error(const xvk_scp* s) { return s ? s->err.c_str() : "null scp"; }
void xvk_scp_close(xvk_scp* s) { delete s; }

}  // extern "C"

// ================================================================================================ writer
struct xvk_writer {
  int fd = -1;
  gzFile gz = nullptr;
  pid_t child = -1;
  std::string cmd, err, ark_path;
  FILE* scp = nullptr;
  std::vector<uint8_t> buf;
  size_t used = 0;
  int64_t offset = 0;       // bytes handed to the stream so far (scp offsets)
};

namespace {

int writer_drain(xvk_writer* w) {
  const uint8_t* p = w->buf.data();
  size_t n = w->used;
  w->used = 0;
  if (w->gz) {
    if (n > 0 && gzwrite(w->gz, p, static_cast<unsigned>(n)) <= 0) { w->err = "gzwrite failed"; return XVK_EIO; }
    return XVK_OK;
  }
  while (n > 0) {
    const ssize_t put = ::write(w->fd, p, n);
    if (put < 0 && errno == EINTR) continue;
    if (put < 0) { set_err(&w->err, "write(): %s", strerror(errno)); return XVK_EIO; }
    p += put;
    n -= static_cast<size_t>(put);
  }
  return XVK_OK;
}

int writer_put(xvk_writer* w, const void* src, size_t n) {
  w->offset += static_cast<int64_t>(n);
  if (n >= w->buf.size()) {                     // large body: flush what is buffered, then write through
    int rc = writer_drain(w);
    if (rc != XVK_OK) return rc;
    const uint8_t* p = static_cast<const uint8_t*>(src);
    if (w->gz) {
      while (n > 0) {
        const unsigned t = n > (1u << 30) ? (1u << 30) : static_cast<unsigned>(n);
        if (gzwrite(w->gz, p, t) <= 0) { w->err = "gzwrite failed"; return XVK_EIO; }
        p += t;
        n -= t;
      }
      return XVK_OK;
    }
    while (n > 0) {
      const ssize_t put = ::write(w->fd, p, n);
      if (put < 0 && errno == EINTR) continue;
      if (put < 0) { set_err(&w->err, "write(): %s", strerror(errno)); return XVK_EIO; }
      p += put;
      n -= static_cast<size_t>(put);
    }
    return XVK_OK;
  }
  if (w->used + n > w->buf.size()) {
    const int rc = writer_drain(w);
    if (rc != XVK_OK) return rc;
  }
  memcpy(w->buf.data() + w->used, src, n);
  w->used += n;
  return XVK_OK;
}

// key + ' ' (ark entries), "\0B", token, dimensions (write_mat / write_vec_flt, dataset/kaldi_io.py:1193-1206, 931-944)
int writer_head(xvk_writer* w, const char* key, const char* token, int32_t d0, int32_t d1, bool matrix) {
  int rc = XVK_OK;
  if (key && key[0]) {
    for (const char* p = key; *p; ++p)
      if (isspace(static_cast<unsigned char>(*p))) { set_err(&w->err, "key '%s' contains whitespace", key); return XVK_EINVAL; }
    rc = writer_put(w, key, strlen(key));
    if (rc == XVK_OK) rc = writer_put(w, " ", 1);
    if (rc != XVK_OK) return rc;
    if (w->scp) fprintf(w->scp, "%s %s:%lld\n", key, w->ark_path.c_str(), (long long)w->offset);
  }
  uint8_t head[16];
  size_t n = 0;
  head[n++] = 0;
  head[n++] = 'B';
  memcpy(head + n, token, 3);
  n += 3;
  head[n++] = 4;
  memcpy(head + n, &d0, 4);
  n += 4;
  if (matrix) {
    head[n++] = 4;
    memcpy(head + n, &d1, 4);
    n += 4;
  }
  return writer_put(w, head, n);
}

template <class T>
int write_matrix(xvk_writer* w, const char* key, const T* m, int32_t rows, int32_t cols, int64_t ld) {
  if (!w || rows < 0 || cols < 0 || ld < cols || (!m && static_cast<int64_t>(rows) * cols > 0)) return XVK_EINVAL;
  int rc = writer_head(w, key, sizeof(T) == 4 ? "FM " : "DM ", rows, cols, true);
  if (rc != XVK_OK) return rc;
  if (ld == cols) return writer_put(w, m, static_cast<size_t>(rows) * cols * sizeof(T));
  for (int64_t r = 0; r < rows && rc == XVK_OK; ++r) rc = writer_put(w, m + r * ld, static_cast<size_t>(cols) * sizeof(T));
  return rc;
}

template <class T>
int write_vector(xvk_writer* w, const char* key, const T* v, int32_t dim) {
  if (!w || dim < 0 || (!v && dim > 0)) return XVK_EINVAL;
  const int rc = writer_head(w, key, sizeof(T) == 4 ? "FV " : "DV ", dim, 0, false);
  if (rc != XVK_OK) return rc;
  return writer_put(w, v, static_cast<size_t>(dim) * sizeof(T));
}

}  // namespace

extern "C" {

xvk_writer* xvk_writer_open(const char* wspecifier, char* err, size_t errcap) {
  if (!wspecifier) { copy_err(err, errcap, "null wspecifier"); return nullptr; }
  std::string prefix, file;
  int64_t offset;
  parse_specifier(wspecifier, &prefix, &file, &offset);
  file = trim(file);
  if (offset >= 0) { file += ":" + std::to_string(offset); }      // a trailing :digits is part of an output name
  if (file.empty()) { copy_err(err, errcap, "empty wspecifier"); return nullptr; }
  xvk_writer* w = new xvk_writer();
  w->buf.resize(kBufBytes);
  std::string scp_path;
  const bool both = prefix.find("ark") == 0 && prefix.find(",scp") != std::string::npos;
  if (both) {
    const size_t comma = file.find(',');
    if (comma == std::string::npos) { copy_err(err, errcap, "ark,scp: needs \"ark_path,scp_path\""); delete w; return nullptr; }
    scp_path = trim(file.substr(comma + 1));
    file = trim(file.substr(0, comma));
  }
  std::string msg;
  if (file[0] == '|') {
    w->cmd = trim(file.substr(1));
    w->child = spawn_shell(w->cmd, false, &w->fd, &msg);
    if (w->child < 0) { copy_err(err, errcap, msg); delete w; return nullptr; }
  } else if (ends_with(file, ".gz")) {
    w->gz = gzopen(file.c_str(), "wb");
    if (!w->gz) { set_err(&msg, "cannot open %s: %s", file.c_str(), strerror(errno)); copy_err(err, errcap, msg); delete w; return nullptr; }
  } else {
    w->fd = ::open(file.c_str(), O_WRONLY | O_CREAT | O_TRUNC | O_CLOEXEC, 0644);
    if (w->fd < 0) { set_err(&msg, "cannot open %s: %s", file.c_str(), strerror(errno)); copy_err(err, errcap, msg); delete w; return nullptr; }
  }
  if (both) {
    if (file[0] == '|' || w->gz) { copy_err(err, errcap, "ark,scp: the ark must be a plain file (offsets)"); xvk_writer_close(w); return nullptr; }
    w->ark_path = file;
    w->scp = fopen(scp_path.c_str(), "w");
    if (!w->scp) { set_err(&msg, "cannot open %s: %s", scp_path.c_str(), strerror(errno)); copy_err(err, errcap, msg); xvk_writer_close(w); return nullptr; }
  }
  return w;
}

int xvk_write_mat_f32(xvk_writer* w, const char* key, const float* m, int32_t rows, int32_t cols, int64_t ld) { return write_matrix(w, key, m, rows, cols, ld); }
int xvk_write_mat_f64(xvk_writer* w, const char* key, const double* m, int32_t rows, int32_t cols, int64_t ld) { return write_matrix(w, key, m, rows, cols, ld); }
int xvk_write_vec_f32(xvk_writer* w, const char* key, const float* v, int32_t dim) { return write_vector(w, key, v, dim); }
int xvk_write_vec_f64(xvk_writer* w, const char* key, const double* v, int32_t dim) { return write_vector(w, key, v, dim); }

int xvk_write_vecs_f32(xvk_writer* w, const char* keys, int64_t n, const float* data, int32_t dim, int64_t ld, const uint8_t* skip) {
  if (!w || !keys || n < 0 || (!data && n > 0) || ld < dim) return XVK_EINVAL;
  const char* k = keys;
  for (int64_t i = 0; i < n; ++i) {
    if (!skip || !skip[i]) {
      const int rc = write_vector(w, k, data + i * ld, dim);
      if (rc != XVK_OK) return rc;
    }
    k += strlen(k) + 1;
  }
  return XVK_OK;
}

int xvk_writer_flush(xvk_writer* w) {
  if (!w) return XVK_EINVAL;
  const int rc = writer_drain(w);
  if (rc == XVK_OK && w->gz) gzflush(w->gz, Z_SYNC_FLUSH);
  if (w->scp) fflush(w->scp);
  return rc;
}

const char* xvk_writer_error(const xvk_writer* w) { return w ? w->err.c_str() : "null writer"; }

int xvk_writer_close(xvk_writer* w) {
  if (!w) return XVK_OK;
  int rc = (w->fd >= 0 || w->gz) ? writer_drain(w) : XVK_OK;
  if (w->gz) { if (gzclose(w->gz) != Z_OK && rc == XVK_OK) rc = XVK_EIO; }
  if (w->fd >= 0 && ::close(w->fd) != 0 && rc == XVK_OK) rc = XVK_EIO;
  if (w->scp) fclose(w->scp);
  if (w->child > 0) {
    const int rc2 = reap(w->child, w->cmd, &w->err);
    if (rc == XVK_OK) rc = rc2;
  }
  delete w;
  return rc;
}

// CRC-32C (Castagnoli), slicing-by-8: the checksum of TensorFlow's tensor-bundle blocks and tensors (tf_checkpoint.py).
uint32_t xvk_crc32c(const void* data, size_t n, uint32_t crc) {
  static uint32_t table[8][256];
  static std::atomic<bool> ready(false);
  if (!ready.load(std::memory_order_acquire)) {
    uint32_t t[8][256];
    for (uint32_t i = 0; i < 256; ++i) {
      uint32_t c = i;
      for (int k = 0; k < 8; ++k) c = (c & 1) ? (c >> 1) ^ 0x82F63B78u : c >> 1;
      t[0][i] = c;
    }
    for (uint32_t i = 0; i < 256; ++i)
      for (int k = 1; k < 8; ++k) t[k][i] = (t[k - 1][i] >> 8) ^ t[0][t[k - 1][i] & 0xFF];
    memcpy(table, t, sizeof(t));                 // identical contents from every thread: benign race
    ready.store(true, std::memory_order_release);
  }
  const uint8_t* p = static_cast<const uint8_t*>(data);
  uint32_t c = ~crc;
  while (n >= 8) {
    uint32_t lo, hi;
    memcpy(&lo, p, 4);
    memcpy(&hi, p + 4, 4);
    lo ^= c;
    c = table[7][lo & 0xFF] ^ table[6][(lo >> 8) & 0xFF] ^ table[5][(lo >> 16) & 0xFF] ^ table[4][lo >> 24] ^
        table[3][hi & 0xFF] ^ table[2][(hi >> 8) & 0xFF] ^ table[1][(hi >> 16) & 0xFF] ^ table[0][hi >> 24];
    p += 8;
    n -= 8;
  }
  while (n--) c = table[0][(c ^ *p++) & 0xFF] ^ (c >> 8);
  return ~c;
}

const char* xvk_version(void) { return "xvkaldi 0.1 (ark/scp FM DM CM FV DV text; gzip; pipes)"; }

}  // extern "C"

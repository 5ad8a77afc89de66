// Episode ingestion: native reader for the reference's on-disk episode format.
//
// The reference stores one episode per file with np.savez_compressed (a ZIP archive of .npy
// members, one per spec name: /root/reference/utils/replay_buffer.py:18-23, save_episode) and
// reads it back in DataLoader worker processes with np.load (:26-29, load_episode; :192-212,
// _try_fetch; :57-75, OfflineReplayBuffer._load).  Here the same files are decoded by C++ threads
// (the Python caller releases the GIL across these calls) straight into pinned staging buffers
// that are then copied into the device-resident columns — no worker processes, no pickling.
//
// Host-only code (zlib raw inflate); lives in the same shared library as the kernels.
#include "common.cuh"
#include <atomic>
#include "fast_inflate.h"
#include <fcntl.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <zlib.h>
#include <algorithm>
#include <memory>
#include <string>
#include <vector>

namespace exb {

struct NpzMember {
  std::string name;        // without the ".npy" suffix
  int method;              // 0 = stored, 8 = deflate
  uint64_t comp_size, raw_size, data_off;
  uint32_t crc;            // CRC-32 of the raw member (header + payload), from the directory
  // parsed from the .npy header
  std::string descr;
  bool fortran;
  std::vector<int64_t> shape;
  uint64_t header_len;     // bytes of the .npy header in the raw stream
  uint64_t payload;        // raw_size - header_len
};

struct NpzFile {
  const uint8_t* base = nullptr;
  size_t size = 0;
  std::vector<NpzMember> members;
};

static inline uint16_t rd16(const uint8_t* p) { return (uint16_t)(p[0] | (p[1] << 8)); }
static inline uint32_t rd32(const uint8_t* p) {
  return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
}
static inline uint64_t rd64(const uint8_t* p) { return (uint64_t)rd32(p) | ((uint64_t)rd32(p + 4) << 32); }

// Inflate (or copy) `want` raw bytes of a member starting at raw offset `skip` into dst.
// `crc` (optional) receives the CRC-32 of everything read, skipped bytes included.
static int member_read(const NpzFile& f, const NpzMember& m, uint64_t skip, uint8_t* dst,
                       uint64_t want, uint32_t* crc = nullptr) {
  if (m.data_off > f.size || m.comp_size > f.size - m.data_off) return EXB_EINVAL;
  const uint8_t* src = f.base + m.data_off;
  uLong c = crc32(0L, Z_NULL, 0);
  auto crc_add = [&](const uint8_t* p, uint64_t n) {
    while (crc && n) {
      const uInt step = n < (1u << 30) ? (uInt)n : (1u << 30);
      c = crc32(c, p, step);
      p += step;
      n -= step;
    }
  };
  if (m.method == 0) {
    if (skip + want > m.comp_size) return EXB_EINVAL;
    memcpy(dst, src + skip, want);
    crc_add(src, skip + want);
    if (crc) *crc = (uint32_t)c;
    return EXB_OK;
  }
  z_stream zs;
  memset(&zs, 0, sizeof(zs));
  if (inflateInit2(&zs, -MAX_WBITS) != Z_OK) return EXB_EINVAL;
  zs.next_in = const_cast<Bytef*>(src);
  zs.avail_in = (uInt)m.comp_size;   // episodes are far below 4 GB compressed per member
  uint8_t scratch[4096];
  int rc = Z_OK;
  while (skip > 0 && rc == Z_OK) {
    const uint64_t n = skip < sizeof(scratch) ? skip : sizeof(scratch);
    zs.next_out = scratch;
    zs.avail_out = (uInt)n;
    rc = inflate(&zs, Z_NO_FLUSH);
    skip -= n - zs.avail_out;
    crc_add(scratch, n - zs.avail_out);
    if (zs.avail_out != 0 && rc != Z_OK) break;
  }
  uint64_t left = want;
  uint8_t* out = dst;
  while (left > 0 && (rc == Z_OK || rc == Z_BUF_ERROR)) {
    const uint64_t n = left < (1u << 30) ? left : (1u << 30);
    zs.next_out = out;
    zs.avail_out = (uInt)n;
    rc = inflate(&zs, Z_NO_FLUSH);
    const uint64_t got = n - zs.avail_out;
    out += got;
    left -= got;
    if (got == 0 && rc != Z_OK) break;
  }
  inflateEnd(&zs);
  crc_add(dst, want - left);
  if (crc) *crc = (uint32_t)c;
  return (left == 0 && skip == 0) ? EXB_OK : EXB_EINVAL;
}

// Whole deflated member -> payload in dst, through the fast decoder (fast_inflate.h).  EXB_OK only
// when the stream decoded cleanly AND its CRC-32 matches the directory's; anything else makes the
// caller take the zlib path.
static int member_read_fast(const NpzFile& f, const NpzMember& m, uint8_t* dst) {
  if (m.method != 8 || m.data_off > f.size || m.comp_size > f.size - m.data_off ||
      m.raw_size > (1ull << 40))
    return EXB_EUNSUPPORTED;
  static thread_local std::unique_ptr<finf::Tables> tables;
  static thread_local std::vector<uint8_t> staging;
  if (!tables) tables.reset(new finf::Tables);
  if (staging.size() < m.raw_size + 8) staging.resize(m.raw_size + 8);
  int rc = EXB_EINVAL;
  if (finf::inflate_all(f.base + m.data_off, m.comp_size, staging.data(), m.raw_size,
                        tables.get()) == 0) {
    uLong c = crc32(0L, Z_NULL, 0);
    for (uint64_t o = 0; o < m.raw_size; o += 1u << 30)
      c = crc32(c, staging.data() + o, (uInt)std::min<uint64_t>(1u << 30, m.raw_size - o));
    if ((uint32_t)c == m.crc) {
      memcpy(dst, staging.data() + m.header_len, m.payload);
      rc = EXB_OK;
    }
  }
  if (staging.size() > (8u << 20)) std::vector<uint8_t>().swap(staging);   // keep small ones only
  return rc;
}

// {'descr': '<f4', 'fortran_order': False, 'shape': (1001, 24), }
static bool parse_npy_header(const uint8_t* h, size_t n, NpzMember* m) {
  if (n < 10 || memcmp(h, "\x93NUMPY", 6) != 0) return false;
  const int major = h[6];
  size_t hlen, off;
  if (major == 1) { hlen = rd16(h + 8); off = 10; }
  else { if (n < 12) return false; hlen = rd32(h + 8); off = 12; }
  if (off + hlen > n) return false;
  const std::string d(reinterpret_cast<const char*>(h + off), hlen);
  m->header_len = off + hlen;
  size_t p = d.find("'descr'");
  if (p == std::string::npos) return false;
  p = d.find('\'', p + 7);
  if (p == std::string::npos) return false;
  size_t q = d.find('\'', p + 1);
  if (q == std::string::npos) return false;
  m->descr = d.substr(p + 1, q - p - 1);
  p = d.find("'fortran_order'");
  if (p == std::string::npos) return false;
  p = d.find_first_not_of(": ", p + 15);
  if (p == std::string::npos) return false;
  m->fortran = d.compare(p, 4, "True") == 0;
  p = d.find("'shape'");
  if (p == std::string::npos) return false;
  p = d.find('(', p);
  q = d.find(')', p);
  if (p == std::string::npos || q == std::string::npos) return false;
  m->shape.clear();
  int64_t v = -1;
  for (size_t i = p + 1; i <= q; ++i) {
    const char c = d[i];
    if (c >= '0' && c <= '9') v = (v < 0 ? 0 : v) * 10 + (c - '0');
    else if (v >= 0) { m->shape.push_back(v); v = -1; }
  }
  return true;
}

static int itemsize_of(const std::string& descr) {
  // '<f4', '|u1', '|b1', '<i8', '<f8', ... (little-endian or byte-order-free scalars only)
  if (descr.size() < 3 || descr[0] == '>') return 0;
  const int n = atoi(descr.c_str() + 2);
  return (n >= 1 && n <= 16) ? n : 0;
}

static int npz_parse(NpzFile* f) {
  // end of central directory: scan back for PK\5\6
  if (f->size < 22) return EXB_EINVAL;
  size_t eocd = std::string::npos;
  const size_t lo = f->size > 65557 ? f->size - 65557 : 0;
  for (size_t i = f->size - 22 + 1; i-- > lo;) {
    if (rd32(f->base + i) == 0x06054b50u) { eocd = i; break; }
  }
  if (eocd == std::string::npos) return EXB_EINVAL;
  uint64_t n_ent = rd16(f->base + eocd + 10), cd_off = rd32(f->base + eocd + 16);
  if (cd_off == 0xffffffffu || n_ent == 0xffffu) {
    // zip64 end-of-central-directory locator sits 20 bytes before the EOCD
    if (eocd < 20 || rd32(f->base + eocd - 20) != 0x07064b50u) return EXB_EINVAL;
    const uint64_t z64 = rd64(f->base + eocd - 20 + 8);
    if (z64 + 56 > f->size || rd32(f->base + z64) != 0x06064b50u) return EXB_EINVAL;
    n_ent = rd64(f->base + z64 + 32);
    cd_off = rd64(f->base + z64 + 48);
  }
  uint64_t p = cd_off;
  for (uint64_t e = 0; e < n_ent; ++e) {
    if (p + 46 > f->size || rd32(f->base + p) != 0x02014b50u) return EXB_EINVAL;
    NpzMember m;
    m.method = rd16(f->base + p + 10);
    m.crc = rd32(f->base + p + 16);
    m.comp_size = rd32(f->base + p + 20);
    m.raw_size = rd32(f->base + p + 24);
    const uint16_t nlen = rd16(f->base + p + 28), xlen = rd16(f->base + p + 30),
                   clen = rd16(f->base + p + 32);
    uint64_t lho = rd32(f->base + p + 42);
    if (p + 46 + nlen + xlen + clen > f->size) return EXB_EINVAL;
    m.name.assign(reinterpret_cast<const char*>(f->base + p + 46), nlen);
    // zip64 extra field (id 1): raw size, compressed size, local header offset, in that order,
    // each present only if the 32-bit field is saturated
    const uint8_t* x = f->base + p + 46 + nlen;
    for (size_t i = 0; i + 4 <= xlen;) {
      const uint16_t id = rd16(x + i), sz = rd16(x + i + 2);
      if (i + 4 + sz > xlen) return EXB_EINVAL;     // extra record runs past the extra field
      if (id == 1) {
        size_t o = i + 4;
        if (m.raw_size == 0xffffffffu && o + 8 <= i + 4 + sz) { m.raw_size = rd64(x + o); o += 8; }
        if (m.comp_size == 0xffffffffu && o + 8 <= i + 4 + sz) { m.comp_size = rd64(x + o); o += 8; }
        if (lho == 0xffffffffu && o + 8 <= i + 4 + sz) { lho = rd64(x + o); }
      }
      i += 4 + sz;
    }
    if (lho > f->size || f->size - lho < 30 || rd32(f->base + lho) != 0x04034b50u)
      return EXB_EINVAL;
    m.data_off = lho + 30 + rd16(f->base + lho + 26) + rd16(f->base + lho + 28);
    // overflow-safe: the member's bytes lie inside the mapping, and a deflate stream expands at
    // most 1032:1, so a directory claiming more than that is damaged (bounds every allocation)
    if (m.data_off > f->size || m.comp_size > f->size - m.data_off) return EXB_EINVAL;
    if (m.raw_size / 1032 > m.comp_size + 1) return EXB_EINVAL;
    if (m.method != 0 && m.method != 8) return EXB_EUNSUPPORTED;
    if (m.name.size() > 4 && m.name.compare(m.name.size() - 4, 4, ".npy") == 0)
      m.name.resize(m.name.size() - 4);
    uint8_t head[1024];
    const uint64_t hn = m.raw_size < sizeof(head) ? m.raw_size : sizeof(head);
    int rc = member_read(*f, m, 0, head, hn);
    if (rc != EXB_OK) return rc;
    if (!parse_npy_header(head, hn, &m) || m.header_len > m.raw_size) return EXB_EINVAL;
    m.payload = m.raw_size - m.header_len;
    f->members.push_back(m);
    p += 46 + nlen + xlen + clen;
  }
  return EXB_OK;
}

}  // namespace exb

using namespace exb;

// No C++ exception may cross the C ABI (std::bad_alloc from a staging buffer, std::length_error /
// out_of_range from header parsing of a damaged file): they become EXB_EINVAL like every other
// unreadable archive, which the loader treats as the reference treats a failed load_episode
// (/root/reference/utils/replay_buffer.py:173-176).
template <typename F>
static int no_throw(const char* what, F&& body) {
  try {
    return body();
  } catch (const std::exception& e) {
    set_error("%s: %s", what, e.what());
  } catch (...) {
    set_error("%s: unknown C++ exception", what);
  }
  return EXB_EINVAL;
}

static int npz_open_impl(const char* path, void** handle, int* n_arrays) {
  EXB_CHECK_ARG(path && handle && n_arrays, "null pointer");
  const int fd = open(path, O_RDONLY);
  if (fd < 0) {
    set_error("cannot open %s", path);
    return EXB_EINVAL;
  }
  struct stat sb;
  if (fstat(fd, &sb) != 0 || sb.st_size <= 0) {
    close(fd);
    set_error("cannot stat %s (or empty file)", path);
    return EXB_EINVAL;
  }
  void* map = mmap(nullptr, (size_t)sb.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
  close(fd);
  if (map == MAP_FAILED) {
    set_error("mmap failed for %s", path);
    return EXB_EINVAL;
  }
  NpzFile* f = new NpzFile;
  f->base = static_cast<const uint8_t*>(map);
  f->size = (size_t)sb.st_size;
  const int rc = npz_parse(f);
  if (rc != EXB_OK) {
    munmap(map, f->size);
    delete f;
    set_error("%s is not a readable .npz archive (status %d)", path, rc);
    return rc;
  }
  *handle = f;
  *n_arrays = (int)f->members.size();
  return EXB_OK;
}

static int npz_info_impl(void* handle, int i, exb_npz_array* out) {
  EXB_CHECK_ARG(handle && out, "null pointer");
  const NpzFile* f = static_cast<const NpzFile*>(handle);
  EXB_CHECK_ARG(i >= 0 && i < (int)f->members.size(), "array index %d out of range", i);
  const NpzMember& m = f->members[i];
  memset(out, 0, sizeof(*out));
  strncpy(out->name, m.name.c_str(), sizeof(out->name) - 1);
  strncpy(out->dtype, m.descr.c_str(), sizeof(out->dtype) - 1);
  out->ndim = (int32_t)(m.shape.size() < 8 ? m.shape.size() : 8);
  for (int d = 0; d < out->ndim; ++d) out->shape[d] = m.shape[d];
  out->nbytes = (int64_t)m.payload;
  out->fortran_order = m.fortran ? 1 : 0;
  const int isz = itemsize_of(m.descr);
  int64_t cnt = 1;
  for (int64_t s : m.shape) cnt *= s;
  if (isz == 0 || m.shape.size() > 8 || cnt * isz != (int64_t)m.payload) {
    set_error("array '%s': unsupported dtype '%s' or inconsistent size", m.name.c_str(),
              m.descr.c_str());
    return EXB_EUNSUPPORTED;
  }
  return EXB_OK;
}

static std::atomic<int> g_inflate_mode{1};

static int npz_read_impl(void* handle, int i, void* host_dst, int64_t dst_bytes) {
  EXB_CHECK_ARG(handle && host_dst, "null pointer");
  const NpzFile* f = static_cast<const NpzFile*>(handle);
  EXB_CHECK_ARG(i >= 0 && i < (int)f->members.size(), "array index %d out of range", i);
  const NpzMember& m = f->members[i];
  EXB_CHECK_ARG(dst_bytes == (int64_t)m.payload, "array '%s': destination has %lld bytes, need %llu",
                m.name.c_str(), (long long)dst_bytes, (unsigned long long)m.payload);
  // 1 (default) fast decoder with zlib behind it, 0 zlib only, 2 fast decoder only (tests: a
  // member it cannot decode is an error instead of a silent slow path): exb_npz_set_inflate_mode
  const int fast_on = g_inflate_mode.load(std::memory_order_relaxed);
  if (fast_on && m.method == 8) {
    if (member_read_fast(*f, m, static_cast<uint8_t*>(host_dst)) == EXB_OK) return EXB_OK;
    if (fast_on == 2) {
      set_error("array '%s': fast decoder failed (strict mode)", m.name.c_str());
      return EXB_EINVAL;
    }
  }
  uint32_t crc = 0;
  const int rc = member_read(*f, m, m.header_len, static_cast<uint8_t*>(host_dst), m.payload, &crc);
  if (rc != EXB_OK) {
    set_error("array '%s': corrupt or truncated stream", m.name.c_str());
    return rc;
  }
  if (crc != m.crc) {   // np.load (zipfile) raises BadZipFile "Bad CRC-32" here
    set_error("array '%s': bad CRC-32 (file %08x, data %08x)", m.name.c_str(), m.crc, crc);
    return EXB_EINVAL;
  }
  return EXB_OK;
}

extern "C" void exb_npz_close(void* handle) {
  if (!handle) return;
  NpzFile* f = static_cast<NpzFile*>(handle);
  munmap(const_cast<uint8_t*>(f->base), f->size);
  delete f;
}

extern "C" int exb_npz_open(const char* path, void** handle, int* n_arrays) {
  return no_throw("exb_npz_open", [&] { return npz_open_impl(path, handle, n_arrays); });
}

extern "C" int exb_npz_info(void* handle, int i, exb_npz_array* out) {
  return no_throw("exb_npz_info", [&] { return npz_info_impl(handle, i, out); });
}

extern "C" int exb_npz_read(void* handle, int i, void* host_dst, int64_t dst_bytes) {
  return no_throw("exb_npz_read", [&] { return npz_read_impl(handle, i, host_dst, dst_bytes); });
}

extern "C" int exb_npz_set_inflate_mode(int mode) {
  const int prev = g_inflate_mode.load(std::memory_order_relaxed);
  if (mode >= 0 && mode <= 2) g_inflate_mode.store(mode, std::memory_order_relaxed);
  return prev;
}

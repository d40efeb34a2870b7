// Gauge-field import straight into the device layout.
//
// QEX's loadGauge (src/gauge/gaugeUtils.nim:87-106) goes through Reader.read
// (src/io/readerQiolite.nim:108-165): the `scidacio` module of the external package qiolite
// (usqcd-software/qiolite#head, qex.nimble:33 -- NOT in the reference tree) hands back this rank's
// hypercube of the record in lexicographic site order, and a CPU loop scatters site by site into the
// SIMD layout, converting precision on the way (put / putP).  Here the raw bytes of the hypercube go
// to the GPU in one copy and ONE kernel byte-swaps, converts float -> double and permutes into the
// checkerboarded SoA link set.  The same kernel imports MILC-order host arrays
// (src/quda/qudaWrapperImpl.nim:216-240: site-major [site][mu][3][3][re,im], native endian).
//
// File format restated from the public SciDAC/LIME conventions (c-lime, QIO): see scidac_* below.
// The reference tree holds no SciDAC file, so this row is pinned only by an independent writer in
// tests/ (format round trip) -- "parity unpinned" against a QEX-written file.
#include <algorithm>
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cstring>

#include "stencil.cuh"

namespace qb {

constexpr int IB = 128;

__device__ __forceinline__ double load_real(const unsigned char* p, int prec, int swap) {
  if (prec == 8) {
    unsigned long long v = *reinterpret_cast<const unsigned long long*>(p);
    if (swap) v = __byte_perm((unsigned)(v >> 32), 0, 0x0123) | ((unsigned long long)__byte_perm((unsigned)v, 0, 0x0123) << 32);
    return __longlong_as_double((long long)v);
  }
  unsigned v = *reinterpret_cast<const unsigned*>(p);
  if (swap) v = __byte_perm(v, 0, 0x0123);
  return (double)__uint_as_float(v);
}

// raw: this rank's sites in LOCAL lexicographic order (x0 fastest), per site nmu matrices of 18 reals
// (row-major, re then im) of `prec` bytes each.  One thread per (site, mu).
__global__ void __launch_bounds__(IB) k_import_sitemajor(double2* __restrict__ u, const unsigned char* __restrict__ raw,
                                                          int nmu, int prec, int swap, int X0, int X1, int X2, int X3,
                                                          int Vh) {
  const long long t = (long long)blockIdx.x * IB + threadIdx.x;
  const long long nthreads = 2ll * Vh * nmu;
  if (t >= nthreads) return;
  const int mu = (int)(t % nmu);
  const int lex = (int)(t / nmu);
  int r = lex;
  const int x0 = r % X0; r /= X0;
  const int x1 = r % X1; r /= X1;
  const int x2 = r % X2;
  const int x3 = r / X2;
  const int par = (x0 + x1 + x2 + x3) & 1, idx = lex >> 1;
  const unsigned char* p = raw + ((size_t)lex * nmu + mu) * 18 * prec;
  double2* o = u + ((size_t)(mu * 2 + par) * 9) * Vh + idx;
#pragma unroll
  for (int e = 0; e < 9; e++)
    o[(size_t)e * Vh] = make_double2(load_real(p + (size_t)(2 * e) * prec, prec, swap),
                                     load_real(p + (size_t)(2 * e + 1) * prec, prec, swap));
}

static unsigned char* g_raw = nullptr;
static size_t g_raw_bytes = 0;
int io_free() {
  if (g_raw) cudaFree(g_raw);
  g_raw = nullptr; g_raw_bytes = 0;
  return 0;
}

int import_sitemajor(Gauge& g, const void* data, int prec, int big_endian) {
  Context& c = ctx();
  if (g.nlinks != 4) return fail("gauge import: the link set must hold 4 link fields");
  if (prec != 4 && prec != 8) return fail("gauge import: precision must be 4 or 8 bytes, got %d", prec);
  const size_t bytes = (size_t)c.g.V * 4 * 18 * prec;
  if (g_raw_bytes < bytes) {
    if (g_raw) QB_CUDA(cudaFree(g_raw));
    g_raw = nullptr; g_raw_bytes = 0;
    QB_CUDA(cudaMalloc(&g_raw, bytes));
    g_raw_bytes = bytes;
  }
  QB_CUDA(cudaMemcpyAsync(g_raw, data, bytes, cudaMemcpyHostToDevice, c.stream));
  const long long nthreads = (long long)c.g.V * 4;
  k_import_sitemajor<<<(unsigned)((nthreads + IB - 1) / IB), IB, 0, c.stream>>>(g.d, g_raw, 4, prec, big_endian ? 1 : 0,
                                                                               c.g.X[0], c.g.X[1], c.g.X[2], c.g.X[3], c.g.Vh);
  count_launch();
  QB_CUDA(cudaGetLastError());
  QB_CUDA(cudaStreamSynchronize(c.stream));
  g.loaded = true;
  return compress_links(g);
}

// ---- SciDAC / LIME ---------------------------------------------------------------------------------
// LIME record: 144-byte header {magic 0x456789ab, version u16, flags u16 (bit 15 MB, bit 14 ME),
// data length u64, type char[128]}, all big-endian, data padded to a multiple of 8 bytes.
// SciDAC lattice file: scidac-private-file-xml (<dims>), scidac-file-xml, then per record
// scidac-private-record-xml (<precision>, <typesize>, <datacount>), scidac-record-xml,
// scidac-binary-data (ILDG: ildg-binary-data) = sites in lexicographic order, x0 fastest, big-endian
// IEEE reals, and scidac-checksum (<suma>, <sumb>): crc32 of every site's bytes, rotated left by
// (site rank mod 29) resp. (mod 31) and XOR-ed together.
static unsigned crc_table[256];
static bool crc_ready = false;
static void crc_init() {
  for (unsigned i = 0; i < 256; i++) {
    unsigned c = i;
    for (int k = 0; k < 8; k++) c = (c & 1) ? 0xedb88320u ^ (c >> 1) : c >> 1;
    crc_table[i] = c;
  }
  crc_ready = true;
}
static unsigned crc32_bytes(const unsigned char* p, size_t n) {
  if (!crc_ready) crc_init();
  unsigned c = 0xffffffffu;
  for (size_t i = 0; i < n; i++) c = crc_table[(c ^ p[i]) & 0xff] ^ (c >> 8);
  return c ^ 0xffffffffu;
}
static inline unsigned rotl32(unsigned v, unsigned r) { return r ? (v << r) | (v >> (32 - r)) : v; }
static unsigned long long be64(const unsigned char* p) {
  unsigned long long v = 0;
  for (int i = 0; i < 8; i++) v = (v << 8) | p[i];
  return v;
}
static bool xml_tag(const std::string& xml, const char* tag, std::string* out) {
  const std::string open = std::string("<") + tag + ">", close = std::string("</") + tag + ">";
  const size_t a = xml.find(open);
  if (a == std::string::npos) return false;
  const size_t b = xml.find(close, a);
  if (b == std::string::npos) return false;
  *out = xml.substr(a + open.size(), b - a - open.size());
  return true;
}

struct ScidacInfo {
  int dims[4] = {0, 0, 0, 0};
  int prec = 0, typesize = 0, datacount = 0;
  long long data_off = -1, data_len = 0;
  unsigned suma = 0, sumb = 0;
  bool have_sum = false;
};

static int scidac_scan(int fd, const char* fn, ScidacInfo* info) {
  struct stat st;
  if (fstat(fd, &st) != 0) return fail("scidac: cannot stat %s", fn);
  long long off = 0;
  while (off + 144 <= (long long)st.st_size) {
    unsigned char h[144];
    if (pread(fd, h, 144, off) != 144) return fail("scidac: short read of a LIME header in %s", fn);
    if (!(h[0] == 0x45 && h[1] == 0x67 && h[2] == 0x89 && h[3] == 0xab)) return fail("scidac: bad LIME magic at offset %lld of %s", off, fn);
    const long long len = (long long)be64(h + 8);
    char type[129]; memcpy(type, h + 16, 128); type[128] = 0;
    const long long data = off + 144;
    if (data + len > (long long)st.st_size) return fail("scidac: record %s overruns the file %s", type, fn);
    const bool is_xml = !strcmp(type, "scidac-private-file-xml") || !strcmp(type, "scidac-private-record-xml") ||
                        !strcmp(type, "scidac-checksum") || !strcmp(type, "ildg-format");
    std::string xml;
    if (is_xml) {
      if (len > (1 << 20)) return fail("scidac: metadata record %s of %lld bytes", type, len);
      xml.resize((size_t)len);
      if (pread(fd, &xml[0], (size_t)len, data) != len) return fail("scidac: short read of %s", type);
    }
    std::string v;
    if (!strcmp(type, "scidac-private-file-xml")) {
      if (!xml_tag(xml, "dims", &v)) return fail("scidac: no <dims> in %s", fn);
      if (sscanf(v.c_str(), "%d %d %d %d", &info->dims[0], &info->dims[1], &info->dims[2], &info->dims[3]) != 4)
        return fail("scidac: <dims> of %s is not four-dimensional: '%s'", fn, v.c_str());
    } else if (!strcmp(type, "scidac-private-record-xml") && info->data_off < 0) {
      if (xml_tag(xml, "precision", &v)) info->prec = (v.find('D') != std::string::npos) ? 8 : 4;
      if (xml_tag(xml, "typesize", &v)) info->typesize = atoi(v.c_str());
      if (xml_tag(xml, "datacount", &v)) info->datacount = atoi(v.c_str());
    } else if ((!strcmp(type, "scidac-binary-data") || !strcmp(type, "ildg-binary-data")) && info->data_off < 0) {
      info->data_off = data; info->data_len = len;
    } else if (!strcmp(type, "scidac-checksum") && info->data_off >= 0 && !info->have_sum) {
      std::string a, b;
      if (xml_tag(xml, "suma", &a) && xml_tag(xml, "sumb", &b)) {
        info->suma = (unsigned)strtoul(a.c_str(), nullptr, 16);
        info->sumb = (unsigned)strtoul(b.c_str(), nullptr, 16);
        info->have_sum = true;
      }
    }
    off = data + ((len + 7) / 8) * 8;
  }
  if (info->dims[0] == 0) return fail("scidac: %s has no scidac-private-file-xml record", fn);
  if (info->data_off < 0) return fail("scidac: %s has no binary data record", fn);
  return 0;
}

// loadGauge(g, fn): first lattice record of the file into the 4 link fields of g.  sums (may be null):
// {file suma, file sumb, this rank's partial suma, partial sumb}; with a single rank a mismatch is an error,
// with several the caller XORs the partials over the ranks and compares.
int read_scidac(Gauge& g, const char* fn, unsigned sums[4]) {
  Context& c = ctx();
  const int fd = open(fn, O_RDONLY);
  if (fd < 0) return fail("scidac: file not found: %s", fn);    // readerQiolite.nim:38-40
  ScidacInfo info;
  int rc = scidac_scan(fd, fn, &info);
  if (rc) { close(fd); return rc; }
  for (int d = 0; d < 4; d++)
    if (info.dims[d] != c.g.lat[d]) {
      close(fd);
      return fail("field lattice [%d %d %d %d] != reader lattice [%d %d %d %d]", c.g.lat[0], c.g.lat[1], c.g.lat[2],
                  c.g.lat[3], info.dims[0], info.dims[1], info.dims[2], info.dims[3]);   // readerQiolite.nim:167-177
    }
  long long gvol = 1;
  for (int d = 0; d < 4; d++) gvol *= c.g.lat[d];
  if (info.prec == 0) info.prec = (info.data_len == gvol * 4 * 18 * 4) ? 4 : 8;   // ILDG files: infer from the size
  if (info.datacount == 0) info.datacount = 4;
  if (info.typesize == 0) info.typesize = 18 * info.prec;
  if (info.datacount != 4) { close(fd); return fail("field count 4 != reader count %d", info.datacount); }   // :181-182
  if (info.typesize != 18 * info.prec) { close(fd); return fail("field data elements 18 != reader data elements %d", info.typesize / info.prec); }
  const size_t site_bytes = (size_t)4 * 18 * info.prec;
  if (info.data_len != (long long)(gvol * site_bytes)) { close(fd); return fail("scidac: binary record holds %lld bytes, lattice needs %lld", info.data_len, (long long)(gvol * site_bytes)); }
  std::vector<unsigned char> buf((size_t)c.g.V * site_bytes);
  const int* X = c.g.X;
  const int* Lg = c.g.lat;
  const size_t row = (size_t)X[0] * site_bytes;
  bool ok = true;
  if (c.g.nranks == 1) {
    // one pread moves at most ~2 GiB on Linux: loop until the record is in (48^3 x 96 fp64 is 6.1 GB)
    size_t done = 0;
    while (done < buf.size()) {
      const size_t want = std::min(buf.size() - done, (size_t)1 << 30);
      const ssize_t n = pread(fd, buf.data() + done, want, info.data_off + (long long)done);
      if (n <= 0) { ok = false; break; }
      done += (size_t)n;
    }
  } else {
    for (int x3 = 0; x3 < X[3] && ok; x3++) for (int x2 = 0; x2 < X[2] && ok; x2++) for (int x1 = 0; x1 < X[1] && ok; x1++) {
      const long long gl = (((long long)(c.g.rc[3] * X[3] + x3) * Lg[2] + (c.g.rc[2] * X[2] + x2)) * Lg[1] + (c.g.rc[1] * X[1] + x1)) * Lg[0] + (long long)c.g.rc[0] * X[0];
      const size_t ll = ((size_t)(x3 * X[2] + x2) * X[1] + x1) * X[0];
      ok = pread(fd, buf.data() + ll * site_bytes, row, info.data_off + gl * (long long)site_bytes) == (ssize_t)row;
    }
  }
  close(fd);
  if (!ok) return fail("scidac: short read of the binary record of %s", fn);
  unsigned suma = 0, sumb = 0;
  for (int x3 = 0; x3 < X[3]; x3++) for (int x2 = 0; x2 < X[2]; x2++) for (int x1 = 0; x1 < X[1]; x1++) for (int x0 = 0; x0 < X[0]; x0++) {
    const unsigned long long gl = (((unsigned long long)(c.g.rc[3] * X[3] + x3) * Lg[2] + (c.g.rc[2] * X[2] + x2)) * Lg[1] + (c.g.rc[1] * X[1] + x1)) * Lg[0] + (c.g.rc[0] * X[0] + x0);
    const size_t ll = ((size_t)(x3 * X[2] + x2) * X[1] + x1) * X[0] + x0;
    const unsigned crc = crc32_bytes(buf.data() + ll * site_bytes, site_bytes);
    suma ^= rotl32(crc, (unsigned)(gl % 29));
    sumb ^= rotl32(crc, (unsigned)(gl % 31));
  }
  if (sums) { sums[0] = info.suma; sums[1] = info.sumb; sums[2] = suma; sums[3] = sumb; }
  if (c.g.nranks == 1 && info.have_sum && (suma != info.suma || sumb != info.sumb))
    return fail("scidac: checksum mismatch in %s: file %08x %08x, data %08x %08x", fn, info.suma, info.sumb, suma, sumb);
  return import_sitemajor(g, buf.data(), info.prec, 1);
}

int scidac_lattice(const char* fn, int dims[4]) {   // getFileLattice, readerQiolite.nim:11-17
  const int fd = open(fn, O_RDONLY);
  if (fd < 0) return fail("scidac: file not found: %s", fn);
  ScidacInfo info;
  const int rc = scidac_scan(fd, fn, &info);
  close(fd);
  if (rc) return rc;
  for (int d = 0; d < 4; d++) dims[d] = info.dims[d];
  return 0;
}

}  // namespace qb

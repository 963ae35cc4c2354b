// exr.cpp — a small OpenEXR reader for environment maps (InputParams::get_exr_image, src/core/loader.rs:374-390, reads the
// first RGBA layer through the `exr` crate): single-part scanline files, HALF / FLOAT channels, compression NONE, RLE, ZIPS
// and ZIP.  Rows come out top to bottom, rgb[(y * w + x) * 3 + c], like the reference's Vec<Vec<Color>>.  Tiled, multi-part,
// deep and PIZ / PXR24 / B44 / DWA files are rejected with a message naming the missing feature.  Twin of spt/exrio.py.
#include <zlib.h>

#include <cstring>
#include <fstream>
#include <map>
#include <sstream>

#include "host.h"

namespace spt_host {
namespace {

float half_to_float(uint16_t h) {
  const uint32_t sign = (uint32_t)(h >> 15) << 31, exp = (h >> 10) & 0x1f, man = h & 0x3ff;
  uint32_t bits;
  if (exp == 0) {
    if (man == 0) bits = sign;
    else { // subnormal: renormalise
      int e = -1; uint32_t m = man;
      do { e++; m <<= 1; } while (!(m & 0x400));
      bits = sign | ((uint32_t)(127 - 15 - e) << 23) | ((m & 0x3ff) << 13);
    }
  } else if (exp == 31) bits = sign | 0x7f800000u | (man << 13);
  else bits = sign | ((exp + 127 - 15) << 23) | (man << 13);
  float f; memcpy(&f, &bits, 4);
  return f;
}

std::string cstr(const std::string& b, size_t& p) {
  const size_t e = b.find('\0', p);
  if (e == std::string::npos) throw HostError("exr: truncated header");
  std::string s = b.substr(p, e - p);
  p = e + 1;
  return s;
}
int32_t rd_i32(const std::string& b, size_t p) { int32_t v; if (p + 4 > b.size()) throw HostError("exr: truncated file"); memcpy(&v, b.data() + p, 4); return v; }
uint64_t rd_u64(const std::string& b, size_t p) { uint64_t v; if (p + 8 > b.size()) throw HostError("exr: truncated file"); memcpy(&v, b.data() + p, 8); return v; }

// inverse of OpenEXR's byte predictor + even/odd interleave
std::string unpredict(std::string t) {
  for (size_t i = 1; i < t.size(); i++) t[i] = (char)((unsigned char)t[i - 1] + (unsigned char)t[i] - 128);
  std::string out(t.size(), '\0');
  const size_t half = (t.size() + 1) / 2;
  for (size_t i = 0; i < half; i++) out[2 * i] = t[i];
  for (size_t i = 0; half + i < t.size(); i++) out[2 * i + 1] = t[half + i];
  return out;
}

std::string unrle(const std::string& src, size_t expected) {
  std::string out;
  size_t p = 0;
  while (p < src.size()) {
    const int c = (signed char)src[p++];
    if (c < 0) { out.append(src, p, (size_t)-c); p += (size_t)-c; }
    else { if (p >= src.size()) break; out.append((size_t)c + 1, src[p++]); }
  }
  if (out.size() != expected) throw HostError("exr: RLE block has the wrong size");
  return out;
}

} // namespace

std::vector<float> load_exr_rgb(const std::string& path, int& w, int& h) {
  std::ifstream fh(path, std::ios::binary);
  if (!fh) throw HostError("can't open '" + path + "'");
  std::ostringstream ss; ss << fh.rdbuf();
  const std::string buf = ss.str();
  if (buf.size() < 8 || (uint32_t)rd_i32(buf, 0) != 20000630u) throw HostError("'" + path + "' is not an OpenEXR file");
  const uint32_t version = (uint32_t)rd_i32(buf, 4);
  if (version & 0x200u) throw HostError("exr: tiled files are not supported");
  if (version & 0x1800u) throw HostError("exr: deep / multi-part files are not supported");
  size_t p = 8;
  std::map<std::string, std::string> attrs;
  while (p < buf.size() && buf[p] != 0) {
    const std::string name = cstr(buf, p);
    cstr(buf, p); // type
    const int32_t size = rd_i32(buf, p); p += 4;
    attrs[name] = buf.substr(p, (size_t)size); p += (size_t)size;
  }
  p++;
  for (const char* need : {"channels", "compression", "dataWindow"}) if (!attrs.count(need)) throw HostError(std::string("exr: header without '") + need + "'");
  const int comp = (unsigned char)attrs["compression"][0];
  static const char* names[] = {"none", "rle", "zips", "zip", "piz", "pxr24", "b44", "b44a", "dwaa", "dwab"};
  if (comp > 3) throw HostError(std::string("exr: compression '") + (comp < 10 ? names[comp] : "?") + "' is not supported (none, rle, zips, zip are)");
  const std::string& dw = attrs["dataWindow"];
  const int x0 = rd_i32(dw, 0), y0 = rd_i32(dw, 4), x1 = rd_i32(dw, 8), y1 = rd_i32(dw, 12);
  w = x1 - x0 + 1; h = y1 - y0 + 1;
  struct Chan { std::string name; int type; };
  std::vector<Chan> chans;
  const std::string& cb = attrs["channels"];
  for (size_t q = 0; q < cb.size() && cb[q] != 0;) {
    Chan c; c.name = cstr(cb, q);
    c.type = rd_i32(cb, q);
    const int xs = rd_i32(cb, q + 8), ys = rd_i32(cb, q + 12);
    q += 16;
    if (xs != 1 || ys != 1) throw HostError("exr: sub-sampled channels are not supported");
    if (c.type != 1 && c.type != 2) throw HostError("exr: UINT channels are not supported");
    chans.push_back(c);
  }
  size_t line_bytes = 0;
  for (const Chan& c : chans) line_bytes += (size_t)w * (c.type == 1 ? 2 : 4);
  const int lpb = comp == 3 ? 16 : 1, n_blocks = (h + lpb - 1) / lpb;
  auto find = [&](const char* n) -> int {
    for (size_t i = 0; i < chans.size(); i++) {
      const std::string& k = chans[i].name;
      if (k == n || (k.size() > strlen(n) + 1 && k.compare(k.size() - strlen(n) - 1, std::string::npos, std::string(".") + n) == 0)) return (int)i;
    }
    return -1;
  };
  int ci[3] = {find("R"), find("G"), find("B")};
  if (ci[0] < 0 || ci[1] < 0 || ci[2] < 0) { const int y = find("Y"); if (y < 0) throw HostError("exr: no R, G, B (or Y) channels"); ci[0] = ci[1] = ci[2] = y; }
  std::vector<float> rgb((size_t)w * h * 3, 0.f);
  for (int b = 0; b < n_blocks; b++) {
    const size_t off = (size_t)rd_u64(buf, p + 8 * (size_t)b);
    const int y = rd_i32(buf, off), size = rd_i32(buf, off + 4);
    if (off + 8 + (size_t)size > buf.size() || y < y0 || y > y1) throw HostError("exr: bad scanline block");
    const int rows = std::min(lpb, y1 - y + 1);
    const size_t expected = (size_t)rows * line_bytes;
    std::string data = buf.substr(off + 8, (size_t)size);
    if (comp != 0 && (size_t)size < expected) { // a block that did not shrink is stored raw
      if (comp == 1) data = unpredict(unrle(data, expected));
      else {
        std::string raw(expected, '\0');
        uLongf n = (uLongf)expected;
        if (uncompress((Bytef*)&raw[0], &n, (const Bytef*)data.data(), (uLong)data.size()) != Z_OK || n != expected) throw HostError("exr: inflate failed");
        data = unpredict(raw);
      }
    }
    if (data.size() != expected) throw HostError("exr: scanline block has the wrong size");
    size_t q = 0;
    for (int r = 0; r < rows; r++) {
      for (size_t c = 0; c < chans.size(); c++) {
        const size_t bpp = chans[c].type == 1 ? 2 : 4;
        for (int k = 0; k < 3; k++) if (ci[k] == (int)c) {
          float* dst = &rgb[((size_t)(y - y0 + r) * w) * 3 + k];
          for (int x = 0; x < w; x++) {
            if (bpp == 2) { uint16_t hv; memcpy(&hv, data.data() + q + 2 * (size_t)x, 2); dst[3 * (size_t)x] = half_to_float(hv); }
            else memcpy(&dst[3 * (size_t)x], data.data() + q + 4 * (size_t)x, 4);
          }
        }
        q += (size_t)w * bpp;
      }
    }
  }
  return rgb;
}

} // namespace spt_host

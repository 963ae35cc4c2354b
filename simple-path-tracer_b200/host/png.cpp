// png.cpp — the image-file ends of the host: 8-bit PNG in (image textures, src/texture/image_tex.rs:155-178 reads
// files through the `image` crate) and 8-bit PNG out (Film::save_image, src/core/film.rs:94-99).  zlib only.
#include <zlib.h>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>

#include "host.h"

namespace spt_host {
namespace {

uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }
void put_be32(std::string& s, uint32_t v) { const char b[4] = {(char)(v >> 24), (char)(v >> 16), (char)(v >> 8), (char)v}; s.append(b, 4); }

void put_chunk(std::string& out, const char* type, const std::string& data) {
  put_be32(out, (uint32_t)data.size());
  std::string body(type, 4); body += data;
  out += body;
  put_be32(out, (uint32_t)crc32(0, (const Bytef*)body.data(), (uInt)body.size()));
}

int paeth(int a, int b, int c) {
  const int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c);
  return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

} // namespace

Image8 load_png(const std::string& path) {
  std::ifstream fh(path, std::ios::binary);
  if (!fh) throw HostError("can't open '" + path + "'");
  std::ostringstream ss; ss << fh.rdbuf();
  const std::string f = ss.str();
  const uint8_t* d = (const uint8_t*)f.data();
  static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
  if (f.size() < 8 || memcmp(d, sig, 8) != 0) throw HostError("not a PNG file");
  int w = 0, h = 0, depth = 0, ctype = 0, interlace = 0;
  std::string idat; std::vector<uint8_t> plte, trns;
  for (size_t p = 8; p + 12 <= f.size();) {
    const uint32_t len = be32(d + p);
    const std::string type(f, p + 4, 4);
    const uint8_t* body = d + p + 8;
    if (p + 12 + len > f.size()) throw HostError("truncated PNG");
    if (type == "IHDR") { w = (int)be32(body); h = (int)be32(body + 4); depth = body[8]; ctype = body[9]; interlace = body[12]; }
    else if (type == "PLTE") plte.assign(body, body + len);
    else if (type == "tRNS") trns.assign(body, body + len);
    else if (type == "IDAT") idat.append((const char*)body, len);
    else if (type == "IEND") break;
    p += 12 + len;
  }
  if (depth != 8 || interlace != 0) throw HostError("only 8-bit non-interlaced PNG is supported");
  const int ch = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : ctype == 6 ? 4 : 0;
  if (!ch || w <= 0 || h <= 0) throw HostError("unsupported PNG colour type");
  const size_t stride = (size_t)w * ch;
  std::vector<uint8_t> raw((stride + 1) * h);
  uLongf rawlen = (uLongf)raw.size();
  if (uncompress(raw.data(), &rawlen, (const Bytef*)idat.data(), (uLong)idat.size()) != Z_OK || rawlen != raw.size()) throw HostError("PNG inflate failed");
  std::vector<uint8_t> px(stride * h);
  for (int y = 0; y < h; y++) {
    const uint8_t* src = &raw[(stride + 1) * y];
    uint8_t* cur = &px[stride * y];
    const uint8_t* up = y ? &px[stride * (y - 1)] : nullptr;
    const int ft = src[0];
    for (size_t i = 0; i < stride; i++) {
      const int a = i >= (size_t)ch ? cur[i - ch] : 0, b = up ? up[i] : 0, c = (up && i >= (size_t)ch) ? up[i - ch] : 0;
      int v = src[1 + i];
      switch (ft) { case 1: v += a; break; case 2: v += b; break; case 3: v += (a + b) / 2; break; case 4: v += paeth(a, b, c); break; default: break; }
      cur[i] = (uint8_t)v;
    }
  }
  Image8 img; img.width = w; img.height = h; img.rgba.resize((size_t)w * h * 4);
  for (size_t i = 0; i < (size_t)w * h; i++) {
    uint8_t r, g, b, a = 255;
    const uint8_t* s = &px[i * ch];
    if (ctype == 0) { r = g = b = s[0]; }
    else if (ctype == 2) { r = s[0]; g = s[1]; b = s[2]; }
    else if (ctype == 3) { const size_t k = s[0]; if (3 * k + 2 >= plte.size()) throw HostError("PNG palette index out of range"); r = plte[3 * k]; g = plte[3 * k + 1]; b = plte[3 * k + 2]; if (k < trns.size()) a = trns[k]; }
    else if (ctype == 4) { r = g = b = s[0]; a = s[1]; }
    else { r = s[0]; g = s[1]; b = s[2]; a = s[3]; }
    img.rgba[4 * i] = r; img.rgba[4 * i + 1] = g; img.rgba[4 * i + 2] = b; img.rgba[4 * i + 3] = a;
  }
  return img;
}

void save_png(const std::string& path, int width, int height, const uint8_t* rgb) {
  std::string raw;
  raw.reserve(((size_t)width * 3 + 1) * height);
  for (int y = 0; y < height; y++) { raw.push_back(0); raw.append((const char*)rgb + (size_t)y * width * 3, (size_t)width * 3); }
  uLongf clen = compressBound((uLong)raw.size());
  std::string comp(clen, '\0');
  if (compress2((Bytef*)&comp[0], &clen, (const Bytef*)raw.data(), (uLong)raw.size(), 6) != Z_OK) throw HostError("PNG deflate failed");
  comp.resize(clen);
  std::string out("\x89PNG\r\n\x1a\n", 8);
  std::string ihdr;
  put_be32(ihdr, (uint32_t)width); put_be32(ihdr, (uint32_t)height);
  ihdr += std::string("\x08\x02\x00\x00\x00", 5); // 8 bit, RGB, deflate, adaptive, no interlace
  put_chunk(out, "IHDR", ihdr);
  put_chunk(out, "IDAT", comp);
  put_chunk(out, "IEND", std::string());
  std::ofstream fh(path, std::ios::binary);
  if (!fh) throw HostError("can't write '" + path + "'");
  fh.write(out.data(), (std::streamsize)out.size());
}

} // namespace spt_host

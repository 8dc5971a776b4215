// asset_io.cpp — the asset path of the raytracing host: Wavefront OBJ/MTL -> the tri_t / tri_ex_t /
// material_info_t arrays the scene builder takes, PNG -> 0x00RRGGBB texels, and a PPM writer.
//
// Reference behaviour (tests/regression/raytracing):
//   mesh.cpp:130-293    load_obj_tiny + Mesh::Mesh: one un-indexed vertex triple per face, texId = the face's material id
//                       (0 when the face has none), one texture per distinct map_Kd name, material colours and scalars
//   surface.cpp:28-55   Surface::loadImage: 8-bit RGB, texel = (R << 16) + (G << 8) + B
//   tracer.cpp:15-33    write_ppm: "P3", rows bottom-up, one "R G B" line per pixel (the text writer that dominates the
//                       reference's wall time) — reproduced byte for byte as format 3; format 6 is the binary equivalent
// The reference parses OBJ/MTL with tinyobjloader 2.0 (vendored header, not part of this repository) and PNG with stb_image.
// Both are restated here from their documented behaviour: file-order faces, '/'-separated 1-based or negative indices,
// quads split along the shorter diagonal, larger polygons ear-clipped in the dominant-axis projection, MTL defaults
// (dissolve 1, shininess 1, ior 1, diffuse 0.6 when only map_Kd is given); PNG = zlib/DEFLATE (RFC 1950/1951) + the five
// scanline filters, non-interlaced.  Checked against the reference's own loader on its shipped assets
// (tests/test_host_assets_cpu.py, through oracle/_ref when present).
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <limits>
#include <sstream>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/vortex_b200_host.h"

#define VXH_EXPORT extern "C" __attribute__((visibility("default")))

namespace {

thread_local std::string g_err;
int fail(const std::string &m) {
  g_err = m;
  return -1;
}

struct MaterialInfo {   // material_info_t, common.h:20-36 (88 bytes)
  float ambient[3], diffuse[3], specular[3], emissive[3];
  float shininess, ior, dissolve, reflectivity;
  int32_t diffuse_tex_id, illum;
  uint32_t tex_width, tex_height;
  uint64_t tex_offset;
};
static_assert(sizeof(MaterialInfo) == 88, "material_info_t layout");

struct Texture {
  std::string name;
  uint32_t w = 0, h = 0;
  std::vector<uint32_t> texels;
};

struct AssetMesh {
  std::vector<float> tris;        // 9 per triangle
  std::vector<uint32_t> triex;    // 16 words per triangle
  std::vector<MaterialInfo> mats;
  std::vector<std::string> tex_names;
  std::vector<uint8_t> tex;       // all textures, concatenated
  uint32_t n_textures = 0;
};

// ------------------------------------------------------------------------------------------------ PNG
struct BitReader {
  const uint8_t *p, *end;
  uint32_t bits = 0;
  int count = 0;
  bool overrun = false;
  uint32_t get(int n) {
    while (count < n) {
      uint32_t b = 0;
      if (p < end) b = *p++;
      else overrun = true;
      bits |= b << count;
      count += 8;
    }
    uint32_t v = bits & ((n == 32) ? 0xFFFFFFFFu : ((1u << n) - 1u));
    bits >>= n;
    count -= n;
    return v;
  }
  void align() { bits = 0, count = 0; }
};

struct Huffman {           // canonical code, decoded bit by bit over per-length counts (RFC 1951 3.2.2)
  uint16_t count[16] = {0}, symbol[288] = {0};
  void build(const uint8_t *lengths, int n) {
    memset(count, 0, sizeof(count));
    for (int i = 0; i < n; i++) count[lengths[i]]++;
    count[0] = 0;
    uint16_t offs[16];
    offs[1] = 0;
    for (int i = 1; i < 15; i++) offs[i + 1] = offs[i] + count[i];
    for (int i = 0; i < n; i++)
      if (lengths[i]) symbol[offs[lengths[i]]++] = (uint16_t)i;
  }
  int decode(BitReader &br) const {
    int code = 0, first = 0, index = 0;
    for (int len = 1; len <= 15; len++) {
      code |= (int)br.get(1);
      int c = count[len];
      if (code - c < first) return symbol[index + (code - first)];
      index += c, first += c;
      first <<= 1, code <<= 1;
      if (br.overrun) return -1;
    }
    return -1;
  }
};

bool inflate_zlib(const std::vector<uint8_t> &in, std::vector<uint8_t> &out) {
  if (in.size() < 6) return false;
  if ((in[0] & 0x0F) != 8 || ((in[0] << 8) | in[1]) % 31 != 0 || (in[1] & 0x20)) return false;
  BitReader br{in.data() + 2, in.data() + in.size()};
  static const uint16_t len_base[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
  static const uint8_t len_extra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
  static const uint16_t dist_base[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
  static const uint8_t dist_extra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};
  static const uint8_t order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
  for (;;) {
    uint32_t last = br.get(1), type = br.get(2);
    if (type == 0) {
      br.align();
      if (br.end - br.p < 4) return false;
      uint32_t len = br.p[0] | (br.p[1] << 8), nlen = br.p[2] | (br.p[3] << 8);
      br.p += 4;
      if ((len ^ 0xFFFF) != nlen || (uint32_t)(br.end - br.p) < len) return false;
      out.insert(out.end(), br.p, br.p + len);
      br.p += len;
    } else if (type == 1 || type == 2) {
      Huffman lit, dist;
      uint8_t lengths[320];
      if (type == 1) {
        for (int i = 0; i < 144; i++) lengths[i] = 8;
        for (int i = 144; i < 256; i++) lengths[i] = 9;
        for (int i = 256; i < 280; i++) lengths[i] = 7;
        for (int i = 280; i < 288; i++) lengths[i] = 8;
        lit.build(lengths, 288);
        for (int i = 0; i < 30; i++) lengths[i] = 5;
        dist.build(lengths, 30);
      } else {
        int hlit = (int)br.get(5) + 257, hdist = (int)br.get(5) + 1, hclen = (int)br.get(4) + 4;
        if (hlit > 286 || hdist > 30) return false;
        uint8_t cl[19] = {0};
        for (int i = 0; i < hclen; i++) cl[order[i]] = (uint8_t)br.get(3);
        Huffman clh;
        clh.build(cl, 19);
        int i = 0;
        while (i < hlit + hdist) {
          int sym = clh.decode(br);
          if (sym < 0) return false;
          if (sym < 16) {
            lengths[i++] = (uint8_t)sym;
          } else {
            int rep, val = 0;
            if (sym == 16) {
              if (i == 0) return false;
              val = lengths[i - 1], rep = 3 + (int)br.get(2);
            } else if (sym == 17) {
              rep = 3 + (int)br.get(3);
            } else {
              rep = 11 + (int)br.get(7);
            }
            if (i + rep > hlit + hdist) return false;
            while (rep--) lengths[i++] = (uint8_t)val;
          }
        }
        lit.build(lengths, hlit);
        dist.build(lengths + hlit, hdist);
      }
      for (;;) {
        int sym = lit.decode(br);
        if (sym < 0 || br.overrun) return false;
        if (sym < 256) {
          out.push_back((uint8_t)sym);
        } else if (sym == 256) {
          break;
        } else {
          sym -= 257;
          if (sym >= 29) return false;
          uint32_t len = len_base[sym] + br.get(len_extra[sym]);
          int ds = dist.decode(br);
          if (ds < 0 || ds >= 30) return false;
          uint32_t d = dist_base[ds] + br.get(dist_extra[ds]);
          if (d > out.size()) return false;
          size_t from = out.size() - d;
          for (uint32_t k = 0; k < len; k++) out.push_back(out[from + k]);
        }
      }
    } else {
      return false;
    }
    if (last) break;
  }
  return !br.overrun;
}

inline uint32_t be32(const uint8_t *p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

// decodes to 8-bit RGB packed as the reference packs it: (R << 16) + (G << 8) + B (surface.cpp:44-49)
int png_decode(const std::vector<uint8_t> &file, uint32_t &w, uint32_t &h, std::vector<uint32_t> &texels) {
  static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
  if (file.size() < 8 || memcmp(file.data(), sig, 8) != 0) return fail("not a PNG file");
  size_t pos = 8;
  int depth = 0, ctype = -1, interlace = 0;
  std::vector<uint8_t> idat, plte;
  w = h = 0;
  while (pos + 8 <= file.size()) {
    uint32_t len = be32(&file[pos]);
    const uint8_t *type = &file[pos + 4];
    if (pos + 12 + (size_t)len > file.size()) return fail("truncated PNG chunk");
    const uint8_t *data = &file[pos + 8];
    if (!memcmp(type, "IHDR", 4)) {
      if (len < 13) return fail("bad IHDR");
      w = be32(data), h = be32(data + 4), depth = data[8], ctype = data[9], interlace = data[12];
    } else if (!memcmp(type, "PLTE", 4)) {
      plte.assign(data, data + len);
    } else if (!memcmp(type, "IDAT", 4)) {
      idat.insert(idat.end(), data, data + len);
    } else if (!memcmp(type, "IEND", 4)) {
      break;
    }
    pos += 12 + (size_t)len;
  }
  if (w == 0 || h == 0 || w > (1u << 24) || h > (1u << 24)) return fail("bad PNG dimensions");
  if (interlace) return fail("interlaced PNG is not supported");
  int channels = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : ctype == 6 ? 4 : 0;
  if (!channels || !(depth == 1 || depth == 2 || depth == 4 || depth == 8 || depth == 16)) return fail("unsupported PNG colour type / depth");
  if (depth < 8 && !(ctype == 0 || ctype == 3)) return fail("unsupported PNG colour type / depth");
  std::vector<uint8_t> raw;
  if (!inflate_zlib(idat, raw)) return fail("corrupt PNG image data");
  const size_t bpp = (size_t)(channels * depth + 7) / 8;            // filter unit
  const size_t stride = ((size_t)w * channels * depth + 7) / 8;
  if (raw.size() < (stride + 1) * h) return fail("short PNG image data");
  std::vector<uint8_t> prev(stride, 0), cur(stride);
  texels.assign((size_t)w * h, 0);
  for (uint32_t y = 0; y < h; y++) {
    const uint8_t *src = &raw[(stride + 1) * y];
    uint8_t ft = src[0];
    for (size_t i = 0; i < stride; i++) {
      int a = i >= bpp ? cur[i - bpp] : 0, b = prev[i], c = i >= bpp ? prev[i - bpp] : 0, x = src[1 + i];
      int v;
      switch (ft) {
        case 0: v = x; break;
        case 1: v = x + a; break;
        case 2: v = x + b; break;
        case 3: v = x + ((a + b) >> 1); break;
        case 4: {
          int p = a + b - c, pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
          v = x + ((pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c));
          break;
        }
        default: return fail("bad PNG filter type");
      }
      cur[i] = (uint8_t)v;
    }
    for (uint32_t x = 0; x < w; x++) {
      uint32_t r, g, bl;
      auto sample = [&](size_t ch) -> uint32_t {   // 8-bit value of channel ch of pixel x
        if (depth == 8) return cur[(size_t)x * channels + ch];
        if (depth == 16) return cur[((size_t)x * channels + ch) * 2];   // high byte, as stb_image reduces 16 -> 8
        size_t bit = (size_t)x * depth;
        uint32_t v = (cur[bit >> 3] >> (8 - depth - (bit & 7))) & ((1u << depth) - 1);
        return v;
      };
      if (ctype == 3) {
        uint32_t idx = sample(0);
        if ((size_t)idx * 3 + 2 < plte.size()) r = plte[idx * 3], g = plte[idx * 3 + 1], bl = plte[idx * 3 + 2];
        else r = g = bl = 0;
      } else if (ctype == 0 || ctype == 4) {
        uint32_t v = sample(0);
        if (depth < 8) v = v * (255u / ((1u << depth) - 1));          // 1/2/4-bit gray scaled to 0..255
        r = g = bl = v;
      } else {
        r = sample(0), g = sample(1), bl = sample(2);
      }
      texels[(size_t)y * w + x] = (r << 16) + (g << 8) + bl;
    }
    prev.swap(cur);
  }
  return 0;
}

bool read_file(const std::string &path, std::vector<uint8_t> &out) {
  FILE *f = fopen(path.c_str(), "rb");
  if (!f) return false;
  fseek(f, 0, SEEK_END);
  long n = ftell(f);
  fseek(f, 0, SEEK_SET);
  out.resize(n > 0 ? (size_t)n : 0);
  size_t got = out.empty() ? 0 : fread(out.data(), 1, out.size(), f);
  fclose(f);
  return got == out.size();
}

}  // namespace

// ------------------------------------------------------------------------------------------------ OBJ / MTL
namespace {

struct VIdx {
  int v, vt, vn;   // 0-based, -1 = absent
};

std::string dir_of(const std::string &path) {
  size_t p = path.find_last_of("/\\");
  return p == std::string::npos ? std::string() : path.substr(0, p);
}

bool fix_index(int idx, int n, int &out) {   // 1-based or negative (relative to the vertices read so far); 0 is invalid
  if (idx > 0) {
    out = idx - 1;
    return true;
  }
  if (idx < 0) {
    out = n + idx;
    return true;
  }
  return false;
}

bool parse_vertex_ref(const char *tok, int nv, int nvt, int nvn, VIdx &r) {
  r.v = r.vt = r.vn = -1;
  char *end = nullptr;
  long a = strtol(tok, &end, 10);
  if (end == tok || !fix_index((int)a, nv, r.v)) return false;
  if (*end != '/') return true;
  ++end;
  if (*end == '/') {   // v//vn
    ++end;
    const char *s = end;
    long c = strtol(s, &end, 10);
    if (end != s && !fix_index((int)c, nvn, r.vn)) return false;
    return true;
  }
  const char *s = end;
  long b = strtol(s, &end, 10);
  if (end != s && !fix_index((int)b, nvt, r.vt)) return false;
  if (*end != '/') return true;
  ++end;
  s = end;
  long c = strtol(s, &end, 10);
  if (end != s && !fix_index((int)c, nvn, r.vn)) return false;
  return true;
}

void material_defaults(MaterialInfo &m) {
  memset(&m, 0, sizeof(m));
  m.dissolve = 1.0f, m.shininess = 1.0f, m.ior = 1.0f;
  m.diffuse_tex_id = -1;
}

// the file name of a map_* statement: options ("-o u v w", "-s ...", "-bm x", "-clamp on", ...) come first
std::string texture_name(std::istringstream &ss) {
  std::string tok, name;
  while (ss >> tok) {
    if (tok[0] == '-' && tok.size() > 1 && !isdigit((unsigned char)tok[1])) {
      int nargs = 1;
      if (tok == "-o" || tok == "-s" || tok == "-t") nargs = 3;
      else if (tok == "-mm") nargs = 2;
      for (int i = 0; i < nargs; i++) {
        std::streampos before = ss.tellg();
        std::string arg;
        if (!(ss >> arg)) break;
        if (nargs == 3 && i > 0) {   // -o/-s/-t take 1 to 3 numbers
          char *e = nullptr;
          strtod(arg.c_str(), &e);
          if (e == arg.c_str()) {
            ss.seekg(before);
            break;
          }
        }
      }
      continue;
    }
    name = tok;
    std::string rest;
    std::getline(ss, rest);   // names with blanks
    while (!rest.empty() && (rest.back() == '\r' || rest.back() == ' ' || rest.back() == '\t')) rest.pop_back();
    name += rest;
    break;
  }
  return name;
}

void load_mtl(const std::string &path, std::vector<MaterialInfo> &mats, std::vector<std::string> &diffuse_names,
              std::unordered_map<std::string, int> &by_name) {
  std::ifstream f(path);
  if (!f.is_open()) return;   // a missing material library is a warning in the reference's loader, not an error
  std::string line;
  int cur = -1;
  bool has_kd = false, has_d = false;
  while (std::getline(f, line)) {
    std::istringstream ss(line);
    std::string key;
    if (!(ss >> key) || key[0] == '#') continue;
    if (key == "newmtl") {
      std::string name;
      std::getline(ss, name);
      size_t b = name.find_first_not_of(" \t");
      name = b == std::string::npos ? std::string() : name.substr(b);
      while (!name.empty() && (name.back() == '\r' || name.back() == ' ' || name.back() == '\t')) name.pop_back();
      MaterialInfo m;
      material_defaults(m);
      mats.push_back(m);
      diffuse_names.emplace_back();
      cur = (int)mats.size() - 1;
      by_name[name] = cur;   // a repeated name maps to the later definition
      has_kd = has_d = false;
      continue;
    }
    if (cur < 0) continue;
    MaterialInfo &m = mats[cur];
    auto read3 = [&](float *dst) {
      float v[3] = {0, 0, 0};
      ss >> v[0] >> v[1] >> v[2];
      dst[0] = v[0], dst[1] = v[1], dst[2] = v[2];
    };
    if (key == "Ka") read3(m.ambient);
    else if (key == "Kd") read3(m.diffuse), has_kd = true;
    else if (key == "Ks") read3(m.specular);
    else if (key == "Ke") read3(m.emissive);
    else if (key == "Ns") ss >> m.shininess;
    else if (key == "Ni") ss >> m.ior;
    else if (key == "illum") ss >> m.illum;
    else if (key == "d") ss >> m.dissolve, has_d = true;
    else if (key == "Tr") {
      float tr = 0;
      ss >> tr;
      if (!has_d) m.dissolve = 1.0f - tr;   // `d` wins over `Tr` when both are present
    } else if (key == "map_Kd") {
      diffuse_names[cur] = texture_name(ss);
      if (!has_kd) m.diffuse[0] = m.diffuse[1] = m.diffuse[2] = 0.6f;
    }
  }
}

// polygons of more than four vertices: ear clipping in the projection on the two axes of the first non-degenerate corner
void ear_clip(const std::vector<VIdx> &face, const std::vector<float> &v, std::vector<VIdx> &out) {
  const size_t n0 = face.size();
  size_t axes[2] = {1, 2};
  for (size_t k = 0; k < n0; ++k) {
    const VIdx &i0 = face[k % n0], &i1 = face[(k + 1) % n0], &i2 = face[(k + 2) % n0];
    const float *p0 = &v[3 * (size_t)i0.v], *p1 = &v[3 * (size_t)i1.v], *p2 = &v[3 * (size_t)i2.v];
    float e0x = p1[0] - p0[0], e0y = p1[1] - p0[1], e0z = p1[2] - p0[2];
    float e1x = p2[0] - p1[0], e1y = p2[1] - p1[1], e1z = p2[2] - p1[2];
    float cx = std::fabs(e0y * e1z - e0z * e1y), cy = std::fabs(e0z * e1x - e0x * e1z), cz = std::fabs(e0x * e1y - e0y * e1x);
    const float eps = std::numeric_limits<float>::epsilon();
    if (cx > eps || cy > eps || cz > eps) {
      if (!(cx > cy && cx > cz)) {
        axes[0] = 0;
        if (cz > cx && cz > cy) axes[1] = 1;
      }
      break;
    }
  }
  std::vector<VIdx> rem = face;
  size_t guess = 0, iterations = n0, prev_n = rem.size();
  while (rem.size() > 3 && iterations > 0) {
    size_t n = rem.size();
    if (guess >= n) guess -= n;
    if (prev_n != n) prev_n = n, iterations = n;
    else iterations--;
    VIdx ind[3];
    float vx[3], vy[3];
    for (size_t k = 0; k < 3; k++) {
      ind[k] = rem[(guess + k) % n];
      vx[k] = v[3 * (size_t)ind[k].v + axes[0]];
      vy[k] = v[3 * (size_t)ind[k].v + axes[1]];
    }
    float e0x = vx[1] - vx[0], e0y = vy[1] - vy[0], e1x = vx[2] - vx[1], e1y = vy[2] - vy[1];
    float cross = e0x * e1y - e0y * e1x;
    float area = (vx[0] * vy[1] - vy[0] * vx[1]) * 0.5f;
    if (cross * area < 0.0f) {   // reflex corner
      guess += 1;
      continue;
    }
    bool overlap = false;
    for (size_t other = 3; other < n; ++other) {
      size_t idx = (guess + other) % n;
      float tx = v[3 * (size_t)rem[idx].v + axes[0]], ty = v[3 * (size_t)rem[idx].v + axes[1]];
      bool inside = false;   // crossing-number test against the candidate ear
      for (int i = 0, j = 2; i < 3; j = i++)
        if (((vy[i] > ty) != (vy[j] > ty)) && (tx < (vx[j] - vx[i]) * (ty - vy[i]) / (vy[j] - vy[i]) + vx[i])) inside = !inside;
      if (inside) {
        overlap = true;
        break;
      }
    }
    if (overlap) {
      guess += 1;
      continue;
    }
    out.push_back(ind[0]), out.push_back(ind[1]), out.push_back(ind[2]);
    rem.erase(rem.begin() + (long)((guess + 1) % n));
  }
  if (rem.size() == 3) out.push_back(rem[0]), out.push_back(rem[1]), out.push_back(rem[2]);
}

int load_obj(const std::string &path, AssetMesh &mesh) {
  std::ifstream f(path);
  if (!f.is_open()) return fail("cannot open " + path);
  const std::string base = dir_of(path);
  std::vector<float> v, vn, vt;
  std::vector<std::string> diffuse_names;
  std::unordered_map<std::string, int> mat_by_name;
  std::vector<VIdx> corners;        // 3 per output triangle
  std::vector<int> tri_material;
  int material = -1;
  std::string line;
  std::vector<VIdx> face;
  while (std::getline(f, line)) {
    const char *p = line.c_str();
    while (*p == ' ' || *p == '\t') ++p;
    if (*p == '\0' || *p == '#' || *p == '\r') continue;
    if (p[0] == 'v' && (p[1] == ' ' || p[1] == '\t')) {
      float x = 0, y = 0, z = 0;
      sscanf(p + 2, "%f %f %f", &x, &y, &z);
      v.push_back(x), v.push_back(y), v.push_back(z);
    } else if (p[0] == 'v' && p[1] == 'n' && (p[2] == ' ' || p[2] == '\t')) {
      float x = 0, y = 0, z = 0;
      sscanf(p + 3, "%f %f %f", &x, &y, &z);
      vn.push_back(x), vn.push_back(y), vn.push_back(z);
    } else if (p[0] == 'v' && p[1] == 't' && (p[2] == ' ' || p[2] == '\t')) {
      float x = 0, y = 0;
      sscanf(p + 3, "%f %f", &x, &y);
      vt.push_back(x), vt.push_back(y);
    } else if (p[0] == 'f' && (p[1] == ' ' || p[1] == '\t')) {
      face.clear();
      std::istringstream ss(p + 2);
      std::string tok;
      bool bad = false;
      while (ss >> tok) {
        VIdx r;
        if (!parse_vertex_ref(tok.c_str(), (int)v.size() / 3, (int)vt.size() / 2, (int)vn.size() / 3, r)) {
          bad = true;
          break;
        }
        face.push_back(r);
      }
      if (bad) return fail("bad face statement in " + path + ": " + line);
      if (face.size() < 3) continue;   // degenerate face: skipped with a warning by the reference's loader
      bool in_range = true;
      for (const VIdx &r : face)
        if (r.v < 0 || (size_t)r.v * 3 + 2 >= v.size()) in_range = false;
      if (!in_range) continue;         // "Face with invalid vertex index found": skipped
      size_t before = corners.size();
      if (face.size() == 3) {
        corners.insert(corners.end(), face.begin(), face.end());
      } else if (face.size() == 4) {   // split along the shorter diagonal
        const float *a = &v[3 * (size_t)face[0].v], *b = &v[3 * (size_t)face[1].v], *c = &v[3 * (size_t)face[2].v], *d = &v[3 * (size_t)face[3].v];
        float e02x = c[0] - a[0], e02y = c[1] - a[1], e02z = c[2] - a[2];
        float e13x = d[0] - b[0], e13y = d[1] - b[1], e13z = d[2] - b[2];
        float s02 = e02x * e02x + e02y * e02y + e02z * e02z, s13 = e13x * e13x + e13y * e13y + e13z * e13z;
        if (s02 < s13) {
          const VIdx t[6] = {face[0], face[1], face[2], face[0], face[2], face[3]};
          corners.insert(corners.end(), t, t + 6);
        } else {
          const VIdx t[6] = {face[0], face[1], face[3], face[1], face[2], face[3]};
          corners.insert(corners.end(), t, t + 6);
        }
      } else {
        ear_clip(face, v, corners);
      }
      for (size_t k = before; k < corners.size(); k += 3) tri_material.push_back(material);
    } else if (!strncmp(p, "usemtl", 6) && (p[6] == ' ' || p[6] == '\t')) {
      std::string name(p + 7);
      size_t b = name.find_first_not_of(" \t");
      name = b == std::string::npos ? std::string() : name.substr(b);
      while (!name.empty() && (name.back() == '\r' || name.back() == ' ' || name.back() == '\t')) name.pop_back();
      auto it = mat_by_name.find(name);
      material = it == mat_by_name.end() ? -1 : it->second;
    } else if (!strncmp(p, "mtllib", 6) && (p[6] == ' ' || p[6] == '\t')) {
      std::istringstream ss(p + 7);
      std::string file;
      while (ss >> file) {   // several libraries may be named; the first that opens is used
        std::string full = base.empty() ? file : base + "/" + file;
        std::ifstream probe(full);
        if (probe.is_open()) {
          load_mtl(full, mesh.mats, diffuse_names, mat_by_name);
          break;
        }
      }
    }
  }
  // Mesh::Mesh (mesh.cpp:219-252): one Surface per distinct diffuse texture name, ids in first-use order over the materials
  std::unordered_map<std::string, int> loaded;
  std::vector<Texture> textures;
  for (size_t i = 0; i < mesh.mats.size(); i++) {
    MaterialInfo &m = mesh.mats[i];
    const std::string &name = diffuse_names[i];
    if (name.empty()) {
      m.diffuse_tex_id = -1, m.tex_width = m.tex_height = 0;
      continue;
    }
    auto it = loaded.find(name);
    if (it == loaded.end()) {
      Texture t;
      t.name = name;
      std::vector<uint8_t> file;
      std::string full = base.empty() ? name : base + "/" + name;
      if (!read_file(full, file)) return fail("texture not found: " + full);
      if (png_decode(file, t.w, t.h, t.texels) != 0) return fail(full + ": " + g_err);
      textures.push_back(std::move(t));
      it = loaded.emplace(name, (int)textures.size() - 1).first;
    }
    m.diffuse_tex_id = it->second;
    m.tex_width = textures[it->second].w, m.tex_height = textures[it->second].h;
  }
  std::vector<uint64_t> tex_off(textures.size());
  for (size_t i = 0; i < textures.size(); i++) {
    tex_off[i] = mesh.tex.size();
    const uint8_t *b = reinterpret_cast<const uint8_t *>(textures[i].texels.data());
    mesh.tex.insert(mesh.tex.end(), b, b + textures[i].texels.size() * 4);
    mesh.tex_names.push_back(textures[i].name);
  }
  for (MaterialInfo &m : mesh.mats) m.tex_offset = m.diffuse_tex_id >= 0 ? tex_off[m.diffuse_tex_id] : 0;
  mesh.n_textures = (uint32_t)textures.size();
  // triangles (mesh.cpp:256-281)
  size_t n = corners.size() / 3;
  mesh.tris.resize(n * 9);
  mesh.triex.assign(n * 16, 0);
  for (size_t i = 0; i < n; i++) {
    float *ex = reinterpret_cast<float *>(&mesh.triex[i * 16]);
    for (int k = 0; k < 3; k++) {
      const VIdx &c = corners[i * 3 + k];
      memcpy(&mesh.tris[i * 9 + 3 * k], &v[3 * (size_t)c.v], 12);
      if (c.vn >= 0 && (size_t)c.vn * 3 + 2 < vn.size()) memcpy(ex + 3 * k, &vn[3 * (size_t)c.vn], 12);
      if (c.vt >= 0 && (size_t)c.vt * 2 + 1 < vt.size()) memcpy(ex + 9 + 2 * k, &vt[2 * (size_t)c.vt], 8);
    }
    mesh.triex[i * 16 + 15] = tri_material[i] >= 0 ? (uint32_t)tri_material[i] : 0u;
  }
  return 0;
}

}  // namespace

// ------------------------------------------------------------------------------------------------ C ABI
VXH_EXPORT const char *vxh_asset_last_error(void) { return g_err.c_str(); }

VXH_EXPORT vxh_asset_mesh_h vxh_obj_load(const char *obj_path) {
  if (!obj_path) {
    fail("null path");
    return nullptr;
  }
  AssetMesh *m = new AssetMesh();
  if (load_obj(obj_path, *m) != 0) {
    delete m;
    return nullptr;
  }
  return m;
}

VXH_EXPORT void vxh_asset_mesh_view(vxh_asset_mesh_h h, vxh_asset_mesh_view_t *v) {
  if (!h || !v) return;
  AssetMesh *m = (AssetMesh *)h;
  v->tris = m->tris.data(), v->triex16 = m->triex.data(), v->tri_count = (uint32_t)(m->tris.size() / 9);
  v->mats = m->mats.data(), v->n_mats = (uint32_t)m->mats.size();
  v->tex = m->tex.data(), v->tex_bytes = m->tex.size(), v->n_textures = m->n_textures;
}

VXH_EXPORT const char *vxh_asset_mesh_texture_name(vxh_asset_mesh_h h, uint32_t index) {
  AssetMesh *m = (AssetMesh *)h;
  return (m && index < m->tex_names.size()) ? m->tex_names[index].c_str() : nullptr;
}

VXH_EXPORT void vxh_asset_mesh_free(vxh_asset_mesh_h h) { delete (AssetMesh *)h; }

VXH_EXPORT int vxh_png_load(const char *path, uint32_t *width, uint32_t *height, uint32_t **texels) {
  if (!path || !width || !height || !texels) return fail("null argument");
  std::vector<uint8_t> file;
  if (!read_file(path, file)) return fail(std::string("File not found: ") + path);
  std::vector<uint32_t> px;
  if (png_decode(file, *width, *height, px) != 0) return -1;
  *texels = (uint32_t *)malloc(px.size() * 4);
  if (!*texels) return fail("out of memory");
  memcpy(*texels, px.data(), px.size() * 4);
  return 0;
}

VXH_EXPORT void vxh_free(void *p) { free(p); }

// pixels: 0x00RRGGBB per pixel, row 0 first, as the kernels write the framebuffer.  flip_vertical = 1 writes the rows
// bottom-up, which is what the reference does (tracer.cpp:24-26).
VXH_EXPORT int vxh_ppm_write(const char *path, const uint32_t *pixels, uint32_t width, uint32_t height, int format, int flip_vertical) {
  if (!path || !pixels || (format != 3 && format != 6)) return fail("bad argument");
  FILE *f = fopen(path, "wb");
  if (!f) return fail(std::string("Failed to open output file: ") + path);
  fprintf(f, "P%d\n%u %u\n255\n", format, width, height);
  std::vector<char> row;
  for (uint32_t y = 0; y < height; ++y) {
    const uint32_t *src = pixels + (size_t)(flip_vertical ? height - 1 - y : y) * width;
    row.clear();
    if (format == 6) {
      row.resize((size_t)width * 3);
      for (uint32_t x = 0; x < width; ++x) {
        row[3 * x + 0] = (char)((src[x] >> 16) & 0xFF), row[3 * x + 1] = (char)((src[x] >> 8) & 0xFF), row[3 * x + 2] = (char)(src[x] & 0xFF);
      }
    } else {
      char tmp[16];
      for (uint32_t x = 0; x < width; ++x) {
        int n = snprintf(tmp, sizeof(tmp), "%u %u %u\n", (src[x] >> 16) & 0xFF, (src[x] >> 8) & 0xFF, src[x] & 0xFF);
        row.insert(row.end(), tmp, tmp + n);
      }
    }
    if (fwrite(row.data(), 1, row.size(), f) != row.size()) {
      fclose(f);
      return fail("short write");
    }
  }
  fclose(f);
  return 0;
}

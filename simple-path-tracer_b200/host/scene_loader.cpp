// scene_loader.cpp — json::load_scene / json::load_renderer of the reference (src/loader/json.rs:19-242) for the in-scope
// component set, flattening into the SptSceneDesc of include/spt.h.  Includes the value-changing quirks of the reference:
//   Q13  homogeneous medium reads sigma_s from the key "sigma_a"          medium/homogeneous.rs:22-23
//   Q23  integer JSON literals are not accepted where floats are expected  core/loader.rs:422-427
//   Q8   environment_light_index = number of lights registered before it   core/scene_resources.rs:160-167
//   Q9   HashMap iteration order -> instances and lights are sorted by name (one valid draw of the reference)
#include <algorithm>
#include <cmath>
#include <cstring>
#include <fstream>
#include <functional>
#include <sstream>

#include "host.h"
#include "json.h"
#include "params.h"

namespace spt_host {

std::vector<float> load_npy_f32(const std::string& path, std::vector<size_t>& shape); // meshio.cpp

namespace {

Json parse_file(const std::string& path, const std::string& env) {
  std::ifstream fh(path, std::ios::binary);
  if (!fh) throw HostError(env + " - External json file not found");
  std::ostringstream ss; ss << fh.rdbuf();
  const std::string text = ss.str();
  try { return JsonParser(text).parse(); } catch (const std::runtime_error& e) { throw HostError(path + ": " + e.what()); }
}

// ---- glam Affine3A arithmetic in f32 with glam's operation order -----------------------------------------------------
struct Affine { float m[3][3]; float t[3]; }; // m[col][row]

void mat3_mul_vec(const float m[3][3], const float v[3], float out[3]) {
  for (int r = 0; r < 3; r++) { const float a = m[0][r] * v[0] + m[1][r] * v[1]; out[r] = a + m[2][r] * v[2]; }
}
Affine affine_mul(const Affine& a, const Affine& b) {
  Affine o;
  for (int c = 0; c < 3; c++) mat3_mul_vec(a.m, b.m[c], o.m[c]);
  float t[3]; mat3_mul_vec(a.m, b.t, t);
  for (int r = 0; r < 3; r++) o.t[r] = t[r] + a.t[r];
  return o;
}
Affine affine_identity() { Affine a{}; for (int i = 0; i < 3; i++) a.m[i][i] = 1.f; return a; }
Affine rot(char axis, float ang) {
  const float s = std::sin(ang), c = std::cos(ang);
  Affine a = affine_identity();
  if (axis == 'x') { a.m[1][1] = c; a.m[1][2] = s; a.m[2][1] = -s; a.m[2][2] = c; }
  else if (axis == 'y') { a.m[0][0] = c; a.m[0][2] = -s; a.m[2][0] = s; a.m[2][2] = c; }
  else { a.m[0][0] = c; a.m[0][1] = s; a.m[1][0] = -s; a.m[1][1] = c; }
  return a;
}
const float PI_F = 3.14159265358979323846f;

// Instance::load, src/primitive/instance.rs:40-57: T * (Rz*Rx*Ry) * S * M, angles in degrees.
void instance_transform(Params& p, float out[12]) {
  Affine trans = affine_identity();
  if (p.contains_key("matrix")) {
    const auto m4 = p.get_matrix("matrix");
    for (int c = 0; c < 3; c++) for (int r = 0; r < 3; r++) trans.m[c][r] = m4[r][c];
    for (int r = 0; r < 3; r++) trans.t[r] = m4[r][3];
  }
  if (p.contains_key("scale")) {
    const auto s = p.get_float3("scale");
    Affine a{}; a.m[0][0] = s[0]; a.m[1][1] = s[1]; a.m[2][2] = s[2];
    trans = affine_mul(a, trans);
  }
  if (p.contains_key("rotate")) {
    const auto r = p.get_float3("rotate");
    const Affine rz = rot('z', r[2] * PI_F / 180.0f), rx = rot('x', r[0] * PI_F / 180.0f), ry = rot('y', r[1] * PI_F / 180.0f);
    trans = affine_mul(affine_mul(affine_mul(rz, rx), ry), trans);
  }
  if (p.contains_key("translate")) {
    const auto t = p.get_float3("translate");
    Affine a = affine_identity(); a.t[0] = t[0]; a.t[1] = t[1]; a.t[2] = t[2];
    trans = affine_mul(a, trans);
  }
  for (int c = 0; c < 3; c++) for (int r = 0; r < 3; r++) out[3 * c + r] = trans.m[c][r];
  for (int r = 0; r < 3; r++) out[9 + r] = trans.t[r];
}

// load_from_value_or_external, json.rs:212-242: string = external file, array = recurse, object = item.
void each(const Json& value, const std::string& base_path, const std::string& env, bool allow_array,
          const std::function<void(const Json&)>& fn) {
  if (value.kind == Json::String) {
    const Json sub = parse_file(join_path(dir_of(base_path), value.s), env);
    each(sub, base_path, env, allow_array, fn);
  } else if (value.kind == Json::Array) {
    if (!allow_array) throw HostError(env + " - Field should not be an array");
    for (const auto& e : value.arr) each(e, base_path, env, true, fn);
  } else fn(value);
}

template <class M, class V>
void add_named(M& names, const char* kind, const std::string& name, V index) {
  if (names.count(name)) throw HostError(std::string("Duplicated ") + kind + " name '" + name + "'");
  names[name] = index;
}
template <class M>
auto lookup(const M& names, const char* kind, const std::string& name) -> decltype(names.begin()->second) {
  auto it = names.find(name);
  if (it == names.end()) throw HostError(std::string("There is no ") + kind + " named '" + name + "'");
  return it->second;
}

SptTextureDesc new_tex(int tag) {
  SptTextureDesc t{};
  t.tag = tag; t.child[0] = t.child[1] = -1;
  for (int i = 0; i < 3; i++) { t.value[i] = 0.f; t.tiling[i] = 1.f; t.offset[i] = 0.f; }
  t.mode = t.wrap = -1; t.width = t.height = 0; t.rgba = nullptr;
  return t;
}

void warn_unused(FlatScene& fs, const Params& p) {
  for (const auto& k : p.unused_keys()) fs.warnings.push_back(p.name + " - unused key '" + k + "'");
}

struct NameTables { // the string tables of src/texture/mod.rs etc.
  static int tex_op(const std::string& ty) {
    if (ty == "add") return SPT_TEX_ADD; if (ty == "sub") return SPT_TEX_SUB; if (ty == "mul") return SPT_TEX_MUL; if (ty == "div") return SPT_TEX_DIV;
    return -1;
  }
  static int tex_mode(const std::string& m) {
    if (m == "texcoords") return SPT_TEXMODE_TEXCOORDS; if (m == "position") return SPT_TEXMODE_POSITION; if (m == "normal") return SPT_TEXMODE_NORMAL;
    if (m == "tangent") return SPT_TEXMODE_TANGENT; if (m == "bitangent") return SPT_TEXMODE_BITANGENT;
    return -1;
  }
  static int tex_wrap(const std::string& w) {
    if (w == "repeat") return SPT_TEXWRAP_REPEAT; if (w == "mirror_repeat") return SPT_TEXWRAP_MIRROR_REPEAT;
    if (w == "clamp") return SPT_TEXWRAP_CLAMP; if (w == "mirror_clamp") return SPT_TEXWRAP_MIRROR_CLAMP;
    return -1;
  }
};

void load_texture(FlatScene& fs, Params& p) { // texture/mod.rs:210-243
  p.name = "texture";
  const std::string ty = p.get_str("type"), name = p.get_str("name");
  p.name = "texture-" + ty + "-" + name;
  SptTextureDesc tex;
  std::vector<uint8_t> pixels;
  if (ty == "scalar") {
    tex = new_tex(SPT_TEX_SCALAR);
    const auto v = p.get_float3("value");
    for (int i = 0; i < 3; i++) tex.value[i] = v[i];
  } else if (ty == "image") {
    const std::string path = p.get_file_path("image_file");
    Image8 img;
    try { img = load_png(path); } catch (const HostError& e) { throw HostError(p.name + " - can't read image '" + path + "' (" + e.what() + ")"); }
    tex = new_tex(SPT_TEX_IMAGE); tex.width = img.width; tex.height = img.height; pixels = std::move(img.rgba);
  } else if (ty == "image_array") { // extension (not in the reference): inline RGBA8 array [H][W][4], for tests without image files
    const Json& a = p.get_any("rgba");
    if (a.kind != Json::Array || a.arr.empty() || a.arr[0].kind != Json::Array) throw HostError(p.name + " - 'rgba' should be array");
    tex = new_tex(SPT_TEX_IMAGE); tex.height = (int)a.arr.size(); tex.width = (int)a.arr[0].arr.size();
    for (const auto& row : a.arr) for (const auto& px : row.arr) for (int c = 0; c < 4; c++) pixels.push_back((uint8_t)px.arr.at(c).i);
  } else if (NameTables::tex_op(ty) >= 0) {
    tex = new_tex(NameTables::tex_op(ty));
    tex.child[0] = lookup(fs.texture_names, "texture", p.get_str("t1"));
    tex.child[1] = lookup(fs.texture_names, "texture", p.get_str("t2"));
  } else throw HostError(p.name + ": unknown type '" + ty + "'");
  fs.textures.push_back(tex); fs.texture_pixels.push_back(std::move(pixels));
  if (p.get_bool_or("is_srgb", false)) {
    SptTextureDesc s = new_tex(SPT_TEX_SRGB); s.child[0] = (int)fs.textures.size() - 1;
    fs.textures.push_back(s); fs.texture_pixels.emplace_back();
  }
  if (!p.unused_keys().empty()) { // TexInputModifier::load_with_tex, input_modifier.rs:49-67
    int mode = -1, wrap = -1;
    if (p.contains_key("mode")) {
      const std::string m = p.get_str("mode");
      mode = NameTables::tex_mode(m);
      if (mode < 0) throw HostError(p.name + " - Unknown texture input mode '" + m + "'");
    }
    if (p.contains_key("wrap")) {
      const std::string w = p.get_str("wrap");
      wrap = NameTables::tex_wrap(w);
      if (wrap < 0) throw HostError(p.name + " - Unknown texture input wrap mode '" + w + "'");
    }
    SptTextureDesc im = new_tex(SPT_TEX_INPUT_MODIFIER);
    if (mode == -1 || mode == SPT_TEXMODE_TEXCOORDS) {
      const auto t2 = p.get_float2_or("tiling", {1.f, 1.f}), o2 = p.get_float2_or("offset", {0.f, 0.f});
      im.tiling[0] = t2[0]; im.tiling[1] = t2[1]; im.tiling[2] = 1.f; im.offset[0] = o2[0]; im.offset[1] = o2[1]; im.offset[2] = 0.f;
    } else {
      const auto t3 = p.get_float3_or("tiling", {1.f, 1.f, 1.f}), o3 = p.get_float3_or("offset", {0.f, 0.f, 0.f});
      for (int i = 0; i < 3; i++) { im.tiling[i] = t3[i]; im.offset[i] = o3[i]; }
    }
    im.child[0] = (int)fs.textures.size() - 1; im.mode = mode; im.wrap = wrap;
    fs.textures.push_back(im); fs.texture_pixels.emplace_back();
  }
  add_named(fs.texture_names, "texture", name, (int32_t)fs.textures.size() - 1);
  warn_unused(fs, p);
}

void roughness(FlatScene& fs, Params& p, int32_t tex[4]) {
  if (p.contains_key("roughness")) { tex[2] = tex[3] = lookup(fs.texture_names, "texture", p.get_str("roughness")); return; }
  tex[2] = lookup(fs.texture_names, "texture", p.get_str("roughness_x"));
  tex[3] = lookup(fs.texture_names, "texture", p.get_str("roughness_y"));
}

void load_material(FlatScene& fs, Params& p) { // material/mod.rs:47-75 and each Material::load
  p.name = "material";
  const std::string ty = p.get_str("type"), name = p.get_str("name");
  p.name = "material-" + ty + "-" + name;
  auto T = [&](const char* key) { return lookup(fs.texture_names, "texture", p.get_str(key)); };
  SptMaterialDesc m{};
  m.tex[0] = m.tex[1] = m.tex[2] = m.tex[3] = -1; m.ior = 1.f; m.rough_chan = 0; m.metal_chan = 0;
  auto ior = [&]() { const float a = p.get_float("int_ior"); return a / p.get_float_or("ext_ior", 1.0f); };
  if (ty == "lambert") { m.tag = SPT_MAT_LAMBERT; m.tex[0] = T("albedo"); }
  else if (ty == "pseudo") { m.tag = SPT_MAT_PSEUDO; }
  else if (ty == "conductor") { m.tag = SPT_MAT_CONDUCTOR; m.tex[0] = T("ior"); m.tex[1] = T("ior_k"); roughness(fs, p, m.tex); }
  else if (ty == "dielectric") { m.tag = SPT_MAT_DIELECTRIC; m.ior = ior(); m.tex[0] = T("reflectance"); m.tex[1] = T("transmittance"); roughness(fs, p, m.tex); }
  else if (ty == "plastic") { m.tag = SPT_MAT_PLASTIC; m.ior = ior(); m.tex[0] = T("albedo"); roughness(fs, p, m.tex); }
  else if (ty == "pbr_metallic") { m.tag = SPT_MAT_PBR_METALLIC; m.tex[0] = T("base_color"); roughness(fs, p, m.tex); m.tex[1] = T("metallic"); }
  else if (ty == "pbr_specular") { m.tag = SPT_MAT_PBR_SPECULAR; m.tex[0] = T("diffuse"); m.tex[1] = T("specular"); roughness(fs, p, m.tex); }
  else if (ty == "subsurface") { m.tag = SPT_MAT_SUBSURFACE; m.ior = ior(); m.tex[0] = T("albedo"); m.tex[1] = T("ld"); roughness(fs, p, m.tex); }
  else if (ty == "pndf_conductor" || ty == "pndf_plastic") throw HostError(p.name + ": P-NDF materials are out of scope of the grafted path");
  else throw HostError(p.name + ": unknown type '" + ty + "'");
  fs.materials.push_back(m);
  add_named(fs.material_names, "material", name, (int32_t)fs.materials.size() - 1);
  warn_unused(fs, p);
}

void load_medium(FlatScene& fs, Params& p) { // medium/mod.rs:41-59, homogeneous.rs:21-27
  p.name = "medium";
  const std::string ty = p.get_str("type"), name = p.get_str("name");
  p.name = "medium-" + ty + "-" + name;
  if (ty != "homogeneous") throw HostError(p.name + ": unknown type '" + ty + "'");
  const auto sa = p.get_float3("sigma_a");
  const auto ss = p.get_float3("sigma_a"); // Q13: the reference reads sigma_s from the key "sigma_a"
  SptMediumDesc m{};
  for (int i = 0; i < 3; i++) { m.sigma_a[i] = sa[i]; m.sigma_s[i] = ss[i]; }
  m.asymmetric = p.get_float("asymmetric");
  fs.mediums.push_back(m);
  add_named(fs.medium_names, "medium", name, (int32_t)fs.mediums.size() - 1);
  warn_unused(fs, p);
}

void load_primitive(FlatScene& fs, Params& p) { // primitive/mod.rs:55-77
  p.name = "primitive";
  const std::string ty = p.get_str("type"), name = p.get_str("name");
  p.name = "primitive-" + ty + "-" + name;
  if (ty == "sphere") { // sphere.rs:41-47
    const auto c = p.get_float3_or("center", {0.f, 0.f, 0.f});
    SptSphereDesc s{}; s.center[0] = c[0]; s.center[1] = c[1]; s.center[2] = c[2]; s.radius = p.get_float("radius");
    fs.spheres.push_back(s);
    add_named(fs.primitive_names, "primitive", name, std::make_pair((int32_t)SPT_PRIM_SPHERE, (int32_t)fs.spheres.size() - 1));
  } else if (ty == "trimesh") { // triangle.rs:57-108
    fs.meshes.push_back(load_mesh(p.get_file_path("obj_file")));
    add_named(fs.primitive_names, "primitive", name, std::make_pair((int32_t)SPT_PRIM_TRIMESH, (int32_t)fs.meshes.size() - 1));
  } else if (ty == "cubic_bezier" || ty == "catmull_clark") {
    // excluded by the north star; loadable-but-skipped so a shared primitives file still parses
    fs.primitive_names[name] = std::make_pair(-1, -1);
    fs.warnings.push_back(p.name + ": primitive type '" + ty + "' is out of scope and was skipped");
    return;
  } else throw HostError(p.name + ": unknown type '" + ty + "'");
  warn_unused(fs, p);
}

void load_surface(FlatScene& fs, Params& p) { // surface.rs:121-164
  p.name = "surface";
  const std::string name = p.get_str("name");
  p.name = "surface-" + name;
  SptSurfaceDesc s{};
  s.material = lookup(fs.material_names, "material", p.get_str("material"));
  s.normal_map = s.emissive_map = s.inside_medium = -1;
  if (p.contains_key("normal_map")) s.normal_map = lookup(fs.texture_names, "texture", p.get_str("normal_map"));
  if (p.contains_key("displacement_map")) lookup(fs.texture_names, "texture", p.get_str("displacement_map")); // loaded but unused by the reference too
  const auto e = p.get_float3_or("emissive", {0.f, 0.f, 0.f});
  for (int i = 0; i < 3; i++) s.emissive[i] = e[i];
  if (p.contains_key("emissive_map")) s.emissive_map = lookup(fs.texture_names, "texture", p.get_str("emissive_map"));
  s.double_sided = p.get_bool_or("double_sided", false) ? 1 : 0;
  if (p.contains_key("inside_medium")) s.inside_medium = lookup(fs.medium_names, "medium", p.get_str("inside_medium"));
  fs.surfaces.push_back(s);
  add_named(fs.surface_names, "surface", name, (int32_t)fs.surfaces.size() - 1);
  warn_unused(fs, p);
}

typedef std::vector<std::pair<std::string, SptInstanceDesc>> PendingInst;
typedef std::vector<std::pair<std::string, SptLightDesc>> PendingLight;

void load_instance(FlatScene& fs, Params& p, PendingInst& pending) { // instance.rs:35-85
  p.name = "instance";
  const std::string name = p.get_str("name");
  p.name = "instance-" + name;
  SptInstanceDesc inst{};
  instance_transform(p, inst.xform);
  if (p.contains_key("surface")) inst.surface = lookup(fs.surface_names, "surface", p.get_str("surface"));
  else {
    SptSurfaceDesc s{};
    s.material = lookup(fs.material_names, "material", p.get_str("material"));
    s.normal_map = s.emissive_map = s.inside_medium = -1;
    fs.surfaces.push_back(s);
    inst.surface = (int32_t)fs.surfaces.size() - 1;
  }
  const auto prim = lookup(fs.primitive_names, "primitive", p.get_str("primitive"));
  if (prim.first < 0) throw HostError(p.name + ": references an out-of-scope primitive");
  inst.prim_kind = prim.first; inst.prim_index = prim.second;
  for (const auto& kv : pending) if (kv.first == name) throw HostError("Duplicated instance name '" + name + "'");
  pending.emplace_back(name, inst);
  warn_unused(fs, p);
}

SptLightDesc blank_light(int tag) { SptLightDesc l{}; l.tag = tag; l.instance = -1; return l; }

void load_light(FlatScene& fs, Params& p, PendingLight& pending) { // light/mod.rs:37-58
  p.name = "light";
  const std::string ty = p.get_str("type"), name = p.get_str("name");
  p.name = "light-" + ty + "-" + name;
  SptLightDesc l = blank_light(-1);
  auto set3 = [&](float* dst, const char* key) { const auto v = p.get_float3(key); dst[0] = v[0]; dst[1] = v[1]; dst[2] = v[2]; };
  if (ty == "directional") { l.tag = SPT_LIGHT_DIRECTIONAL; set3(l.direction, "direction"); set3(l.strength, "strength"); }
  else if (ty == "point") { l.tag = SPT_LIGHT_POINT; set3(l.position, "position"); set3(l.strength, "strength"); }
  else if (ty == "spot") { // spot.rs:32-48
    l.tag = SPT_LIGHT_SPOT; set3(l.position, "position"); set3(l.direction, "direction");
    l.inner_angle = p.get_float_or("inner_angle", 0.0f) * PI_F / 180.0f;
    l.outer_angle = p.get_float_or("outer_angle", 90.0f) * PI_F / 180.0f;
    set3(l.strength, "strength");
  } else throw HostError(p.name + ": unknown type '" + ty + "'");
  for (const auto& kv : pending) if (kv.first == name) throw HostError("Duplicated light name '" + name + "'");
  pending.emplace_back(name, l);
  warn_unused(fs, p);
}

// Fixed-formula HDR lat-long map: sky gradient + a sun disc `sun_gain` times brighter.
std::vector<float> procedural_sky(int width, int height, double sun_gain) {
  const double sun_theta = 0.9, sun_phi = 1.0, sun_radius = 0.05, pi = 3.14159265358979323846;
  const double sd[3] = {std::sin(sun_theta) * std::sin(sun_phi), std::cos(sun_theta), std::sin(sun_theta) * std::cos(sun_phi)};
  const double A[3] = {0.9, 0.9, 1.0}, B[3] = {0.25, 0.45, 1.0}, G[3] = {0.22, 0.2, 0.18}, S[3] = {1.0, 0.95, 0.85};
  std::vector<float> out((size_t)width * height * 3);
  for (int y = 0; y < height; y++) {
    const double th = (y + 0.5) / height * pi;
    for (int x = 0; x < width; x++) {
      const double ph = (x + 0.5) / width * 2 * pi;
      const double d[3] = {std::sin(th) * std::sin(ph), std::cos(th), std::sin(th) * std::cos(ph)};
      const double up = std::min(1.0, std::max(-1.0, d[1]));
      const double dot = (d[0] * sd[0] + d[1] * sd[1]) + d[2] * sd[2];
      const double ang = std::acos(std::min(1.0, std::max(-1.0, dot)));
      const double sun = ang < sun_radius ? sun_gain : 0.0;
      for (int c = 0; c < 3; c++) {
        const double sky = up > 0 ? (1 - up) * A[c] + up * B[c] : G[c] * (1 + up * 0.5);
        out[((size_t)y * width + x) * 3 + c] = (float)(sky + sun * S[c]);
      }
    }
  }
  return out;
}

void load_environment(FlatScene& fs, Params& p) { // environment.rs:86-108
  const std::string ty = p.get_str("type");
  const auto scale = p.get_float3_or("scale", {1.f, 1.f, 1.f});
  int w = 1, h = 1;
  std::vector<float> rgb;
  if (ty == "color") { const auto c = p.get_float3("color"); rgb = {c[0], c[1], c[2]}; }
  else if (ty == "exr") { // environment.rs:95-98, core/loader.rs:374-390
    const std::string path = p.get_file_path("exr_file");
    try { rgb = load_exr_rgb(path, w, h); } catch (const HostError& e) { throw HostError(std::string("environment - can't read '") + path + "': " + e.what()); }
  }
  else if (ty == "npy") { // extension: lat-long f32 array [H,W,3]
    std::vector<size_t> shape;
    rgb = load_npy_f32(p.get_file_path("npy_file"), shape);
    if (shape.size() != 3 || shape[2] != 3) throw HostError("environment - 'npy_file' should hold an [H,W,3] array");
    h = (int)shape[0]; w = (int)shape[1];
  } else if (ty == "procedural_sky") { // extension: see procedural_sky()
    w = p.contains_key("width") ? (int)p.get_int("width") : 1024;
    h = p.contains_key("height") ? (int)p.get_int("height") : 512;
    rgb = procedural_sky(w, h, (double)p.get_float_or("sun_gain", 1000.0f));
  } else throw HostError(p.name + " - unknown type");
  if (fs.has_env) throw HostError("Environment has been set before");
  fs.has_env = true;
  fs.env_rgb = std::move(rgb);
  fs.env.width = w; fs.env.height = h;
  for (int i = 0; i < 3; i++) fs.env.scale[i] = scale[i];
  for (const auto& k : p.unused_keys()) fs.warnings.push_back("environment - unused key '" + k + "'");
}

} // namespace

int32_t FlatScene::camera_index(const std::string* name) const { // Scene::get_camera, src/core/scene.rs:29-41
  if (name) {
    auto it = camera_names.find(*name);
    if (it == camera_names.end()) throw HostError("There is no camera names " + *name);
    return it->second;
  }
  if (cameras.size() == 1) return 0;
  throw HostError("There are multiple cameras so a name must be given");
}

const SptSceneDesc& FlatScene::desc() {
  mesh_descs_.clear();
  for (const Mesh& m : meshes) {
    SptMeshDesc d{};
    d.n_verts = (int32_t)(m.positions.size() / 3); d.n_tris = (int32_t)(m.indices.size() / 3);
    d.position = m.positions.data(); d.normal = m.normals.data(); d.texcoord = m.texcoords.data();
    d.tangent = m.tangents.data(); d.bitangent = m.bitangents.data(); d.index = m.indices.data();
    mesh_descs_.push_back(d);
  }
  for (size_t i = 0; i < textures.size(); i++) textures[i].rgba = texture_pixels[i].empty() ? nullptr : texture_pixels[i].data();
  SptSceneDesc& d = desc_;
  d = SptSceneDesc{};
  d.n_meshes = (int32_t)mesh_descs_.size(); d.meshes = mesh_descs_.data();
  d.n_spheres = (int32_t)spheres.size(); d.spheres = spheres.data();
  d.n_instances = (int32_t)instances.size(); d.instances = instances.data();
  d.n_surfaces = (int32_t)surfaces.size(); d.surfaces = surfaces.data();
  d.n_materials = (int32_t)materials.size(); d.materials = materials.data();
  d.n_textures = (int32_t)textures.size(); d.textures = textures.data();
  d.n_mediums = (int32_t)mediums.size(); d.mediums = mediums.data();
  d.n_lights = (int32_t)lights.size(); d.lights = lights.data();
  d.n_cameras = (int32_t)cameras.size(); d.cameras = cameras.data();
  if (has_env) { env.rgb = env_rgb.data(); d.env = &env; }
  d.env_light_index = env_light_index;
  d.light_sampler = light_sampler;
  d.aggregate = aggregate;
  return d;
}

// json::load_scene, src/loader/json.rs:53-199.  Relative files (external JSON, obj_file, image_file) resolve against
// `base_path` (defaults to the scene file).
FlatScene load_scene(const std::string& scene_path, const std::string& base_path_in) {
  const std::string base_path = base_path_in.empty() ? scene_path : base_path_in;
  const Json root = parse_file(scene_path, "json");
  if (root.kind != Json::Object) throw HostError("top - expected an object");
  FlatScene fs;
  auto section = [&](const char* key, const char* env, const std::function<void(Params&)>& fn, bool allow_array, bool required) {
    const Json* v = root.find(key);
    if (!v) { if (required) throw HostError(std::string("top - There is no '") + key + "' field"); return; }
    each(*v, base_path, env, allow_array, [&](const Json& obj) { Params p(obj, env, base_path); fn(p); });
  };
  auto load_camera = [&](Params& p) { // camera/mod.rs:30-48, perspective.rs:29-38
    p.name = "camera";
    const std::string ty = p.get_str("type"), name = p.get_str("name");
    if (ty != "perspective") throw HostError("camera-" + ty + "-" + name + ": unknown type '" + ty + "'");
    SptCameraDesc c{};
    c.fov = p.get_float("fov") * PI_F / 180.0f;
    const auto e = p.get_float3("eye"), f = p.get_float3("forward"), u = p.get_float3("up");
    for (int i = 0; i < 3; i++) { c.eye[i] = e[i]; c.forward[i] = f[i]; c.up[i] = u[i]; }
    fs.cameras.push_back(c);
    add_named(fs.camera_names, "camera", name, (int32_t)fs.cameras.size() - 1);
  };
  PendingInst pend_inst; PendingLight pend_light;
  section("cameras", "json-cameras", load_camera, true, true);
  section("textures", "json-textures", [&](Params& p) { load_texture(fs, p); }, true, true);
  section("materials", "json-materials", [&](Params& p) { load_material(fs, p); }, true, true);
  section("mediums", "json-mediums", [&](Params& p) { load_medium(fs, p); }, true, true);
  section("primitives", "json-primitives", [&](Params& p) { load_primitive(fs, p); }, true, true);
  section("surfaces", "json-surfaces", [&](Params& p) { load_surface(fs, p); }, true, true);
  section("instances", "json-instances", [&](Params& p) { load_instance(fs, p, pend_inst); }, true, true);
  section("lights", "json-lights", [&](Params& p) { load_light(fs, p, pend_light); }, true, true);
  const size_t n_explicit = pend_light.size();
  section("environment", "json-environment", [&](Params& p) { load_environment(fs, p); }, false, false);
  if (fs.has_env) { // SceneResources::add_environment, scene_resources.rs:160-167 (Q8)
    pend_light.emplace_back("$env", blank_light(SPT_LIGHT_ENV));
    fs.env_light_index = (int32_t)n_explicit;
  }
  if (root.find("gltf")) throw HostError("json - 'gltf' import is out of scope of the grafted path");
  if (const Json* a = root.find("aggregate")) {
    if (a->kind == Json::String && a->s == "group") fs.aggregate = SPT_AGG_GROUP;
    else if (a->kind == Json::String && a->s == "bvh") fs.aggregate = SPT_AGG_BVH;
    else throw HostError("Unknown aggregate type '" + a->s + "'");
  }
  if (const Json* l = root.find("light_sampler")) {
    if (l->kind == Json::String && l->s == "uniform") fs.light_sampler = SPT_LS_UNIFORM;
    else if (l->kind == Json::String && l->s == "power_is") fs.light_sampler = SPT_LS_POWER_IS;
    else throw HostError("Unknown light sampler type '" + l->s + "'");
  }
  // Q9: fix the HashMap order by name
  std::stable_sort(pend_inst.begin(), pend_inst.end(), [](const PendingInst::value_type& a, const PendingInst::value_type& b) { return a.first < b.first; });
  std::stable_sort(pend_light.begin(), pend_light.end(), [](const PendingLight::value_type& a, const PendingLight::value_type& b) { return a.first < b.first; });
  for (const auto& kv : pend_inst) { fs.instances.push_back(kv.second); add_named(fs.instance_names, "instance", kv.first, (int32_t)fs.instances.size() - 1); }
  for (const auto& kv : pend_light) { fs.lights.push_back(kv.second); add_named(fs.light_names, "light", kv.first, (int32_t)fs.lights.size() - 1); }
  // build_light_sampler, scene_resources.rs:105-138: one ShapeLight per emissive instance
  for (size_t i = 0; i < fs.instances.size(); i++) {
    const float* e = fs.surfaces[fs.instances[i].surface].emissive;
    const float lum = (0.299f * e[0] + 0.587f * e[1]) + 0.114f * e[2];
    if (lum > 0.f) { SptLightDesc l = blank_light(SPT_LIGHT_SHAPE); l.instance = (int32_t)i; fs.lights.push_back(l); }
  }
  if (fs.cameras.empty()) throw HostError("At least one camera is needed");
  return fs;
}

// json::load_renderer, src/loader/json.rs:19-51.
RendererConfig load_renderer_config(const std::string& path) {
  const Json root = parse_file(path, "renderer");
  if (root.kind != Json::Object) throw HostError("renderer - expected an object");
  for (const char* k : {"max_depth", "sampler", "filter", "type"}) if (!root.find(k)) throw HostError(std::string("renderer - There is no '") + k + "' field");
  if (!root.find("max_depth")->is_int()) throw HostError("renderer - 'max_depth' shoule be integer");
  RendererConfig rc;
  rc.max_depth = (int32_t)root.find("max_depth")->i;
  Params sp(*root.find("sampler"), "sampler", "");
  const std::string ty = sp.get_str("type");
  if (ty == "random" || ty == "recurrence") { rc.sampler.kind = ty == "random" ? SPT_SAMPLER_RANDOM : SPT_SAMPLER_RECURRENCE; rc.sampler.spp = (int32_t)sp.get_int("spp"); }
  else if (ty == "jittered") {
    rc.sampler.kind = SPT_SAMPLER_JITTERED;
    rc.sampler.division_x = (int32_t)sp.get_int("division_x"); rc.sampler.division_y = (int32_t)sp.get_int("division_y");
    rc.sampler.spp = rc.sampler.division_x * rc.sampler.division_y;
  } else throw HostError("sampler-" + ty + ": unknown type '" + ty + "'");
  Params fp(*root.find("filter"), "filter", "");
  const std::string fty = fp.get_str("type");
  if (fty != "box") throw HostError("filter-" + fty + ": unknown type '" + fty + "'");
  rc.filter.radius = fp.get_float("radius");
  if (root.find("type")->kind != Json::String) throw HostError("renderer - 'type' shoule be string");
  rc.type = root.find("type")->s;
  return rc;
}

} // namespace spt_host

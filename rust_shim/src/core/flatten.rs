// src/core/flatten.rs — SceneResources -> FlatScene (the owner of every array an SptSceneDesc points into).
// UNCOMPILED HERE (no Rust toolchain in the build image).  The C++ twin that IS built and tested is
// simple-path-tracer_b200/host/scene_loader.cpp (FlatScene::desc) and the Python one spt/loader.py (FlatScene.to_desc).
//
// Hook: SceneResources::to_scene (scene_resources.rs:28-47) calls `let flat = self.flatten(aggregate_type, light_sampler_type)`
// BEFORE it moves its maps into Scene::new, and Scene gets one more field `flat: Arc<FlatScene>` with
// `pub fn flatten(&self) -> Arc<FlatScene>`.  The private fields read below need `pub(crate)` accessors of the same name
// on Instance (instance.rs:10-16), Surface (surface.rs:14-22), TriMesh/Triangle (triangle.rs:19-27), Sphere
// (sphere.rs:8-12), each Material / Texture / Light variant, Homogeneous (homogeneous.rs:5-9) and PerspectiveCamera
// (perspective.rs:5-12).
use std::{collections::HashMap, sync::Arc};

use crate::{
    core::scene_resources::SceneResources,
    light::Light,
    material::Material,
    primitive::Primitive,
    renderer::ffi::*,
    texture::Texture,
};

#[derive(Default)]
pub struct FlatMesh {
    pub position: Vec<f32>,
    pub normal: Vec<f32>,
    pub texcoord: Vec<f32>,
    pub tangent: Vec<f32>,
    pub bitangent: Vec<f32>,
    pub index: Vec<u32>,
}

#[derive(Default)]
pub struct FlatScene {
    pub meshes: Vec<FlatMesh>,
    pub spheres: Vec<SptSphereDesc>,
    pub instances: Vec<SptInstanceDesc>,
    pub surfaces: Vec<SptSurfaceDesc>,
    pub materials: Vec<SptMaterialDesc>,
    pub textures: Vec<SptTextureDesc>,
    pub texture_pixels: Vec<Vec<u8>>, // parallel to textures; level-0 RGBA8 of ImageTex, else empty
    pub mediums: Vec<SptMediumDesc>,
    pub lights: Vec<SptLightDesc>,
    pub cameras: Vec<SptCameraDesc>,
    pub camera_names: HashMap<String, i32>,
    pub env_rgb: Vec<f32>,
    pub env: Option<SptEnvDesc>,
    pub env_light_index: i32,
    pub light_sampler: i32,
    pub aggregate: i32,
    mesh_descs: Vec<SptMeshDesc>,
}

impl FlatScene {
    /// Scene::get_camera (scene.rs:29-41)
    pub fn camera_index(&self, name: Option<&str>) -> i32 {
        match name {
            Some(n) => *self.camera_names.get(n).unwrap_or_else(|| panic!("There is no camera names {}", n)),
            None if self.cameras.len() == 1 => 0,
            None => panic!("There are multiple cameras so a name must be given"),
        }
    }

    /// Pointers stay valid while `self` is alive and unchanged.
    pub fn desc(&self) -> SptSceneDesc {
        SptSceneDesc {
            n_meshes: self.mesh_descs.len() as i32, meshes: self.mesh_descs.as_ptr(),
            n_spheres: self.spheres.len() as i32, spheres: self.spheres.as_ptr(),
            n_instances: self.instances.len() as i32, instances: self.instances.as_ptr(),
            n_surfaces: self.surfaces.len() as i32, surfaces: self.surfaces.as_ptr(),
            n_materials: self.materials.len() as i32, materials: self.materials.as_ptr(),
            n_textures: self.textures.len() as i32, textures: self.textures.as_ptr(),
            n_mediums: self.mediums.len() as i32, mediums: self.mediums.as_ptr(),
            n_lights: self.lights.len() as i32, lights: self.lights.as_ptr(),
            n_cameras: self.cameras.len() as i32, cameras: self.cameras.as_ptr(),
            env: self.env.as_ref().map_or(std::ptr::null(), |e| e as *const _),
            env_light_index: self.env_light_index,
            light_sampler: self.light_sampler,
            aggregate: self.aggregate,
        }
    }

    fn seal(&mut self) {
        self.mesh_descs = self.meshes.iter().map(|m| SptMeshDesc {
            n_verts: (m.position.len() / 3) as i32, n_tris: (m.index.len() / 3) as i32,
            position: m.position.as_ptr(), normal: m.normal.as_ptr(), texcoord: m.texcoord.as_ptr(),
            tangent: m.tangent.as_ptr(), bitangent: m.bitangent.as_ptr(), index: m.index.as_ptr(),
        }).collect();
        for (t, px) in self.textures.iter_mut().zip(&self.texture_pixels) {
            t.rgba = if px.is_empty() { std::ptr::null() } else { px.as_ptr() };
        }
        if let Some(e) = self.env.as_mut() { e.rgb = self.env_rgb.as_ptr(); }
    }
}

/// Arc identity -> flat index (the reference links components by Arc, the ABI by index)
struct Interner<T> { ids: HashMap<*const T, i32> }
impl<T> Interner<T> {
    fn new() -> Self { Self { ids: HashMap::new() } }
    fn get(&self, a: &Arc<T>) -> Option<i32> { self.ids.get(&Arc::as_ptr(a)).copied() }
    fn put(&mut self, a: &Arc<T>, i: i32) { self.ids.insert(Arc::as_ptr(a), i); }
}

impl SceneResources {
    pub fn flatten(&self, aggregate_type: Option<&str>, light_sampler_type: Option<&str>) -> FlatScene {
        let mut fs = FlatScene::default();
        let mut tex_ids = Interner::<Texture>::new();
        let mut mat_ids = Interner::<Material>::new();
        let mut prim_ids: HashMap<*const Primitive, (i32, i32)> = HashMap::new();

        // textures: children before parents (texture/mod.rs:198-208)
        fn add_tex(fs: &mut FlatScene, ids: &mut Interner<Texture>, t: &Arc<Texture>) -> i32 {
            if let Some(i) = ids.get(t) { return i; }
            let mut d = SptTextureDesc { tag: 0, child: [-1, -1], value: [0.0; 3], tiling: [1.0; 3], offset: [0.0; 3],
                                         mode: -1, wrap: -1, width: 0, height: 0, rgba: std::ptr::null() };
            let mut px = vec![];
            match &**t {
                Texture::ScalarTex(s) => { d.tag = SPT_TEX_SCALAR; d.value = s.value().into(); }
                Texture::ImageTex(im) => { d.tag = SPT_TEX_IMAGE; d.width = im.width() as i32; d.height = im.height() as i32; px = im.level0_rgba8(); }
                Texture::AddTex(b) => { d.tag = SPT_TEX_ADD; d.child = [add_tex(fs, ids, b.t1()), add_tex(fs, ids, b.t2())]; }
                Texture::SubTex(b) => { d.tag = SPT_TEX_SUB; d.child = [add_tex(fs, ids, b.t1()), add_tex(fs, ids, b.t2())]; }
                Texture::MulTex(b) => { d.tag = SPT_TEX_MUL; d.child = [add_tex(fs, ids, b.t1()), add_tex(fs, ids, b.t2())]; }
                Texture::DivTex(b) => { d.tag = SPT_TEX_DIV; d.child = [add_tex(fs, ids, b.t1()), add_tex(fs, ids, b.t2())]; }
                Texture::SrgbTex(s) => { d.tag = SPT_TEX_SRGB; d.child[0] = add_tex(fs, ids, s.inner()); }
                Texture::TexInputModifier(m) => {
                    d.tag = SPT_TEX_INPUT_MODIFIER; d.child[0] = add_tex(fs, ids, m.inner());
                    d.tiling = m.tiling().into(); d.offset = m.offset().into();
                    d.mode = m.mode().map_or(-1, |x| x as i32); d.wrap = m.wrap().map_or(-1, |x| x as i32);
                }
            }
            fs.textures.push(d); fs.texture_pixels.push(px);
            let i = fs.textures.len() as i32 - 1;
            ids.put(t, i);
            i
        }
        for t in self.textures.values() { add_tex(&mut fs, &mut tex_ids, t); }

        // materials (material/mod.rs:33-45); roughness textures in tex[2..3], channel selectors for the pbr_* variants
        for m in self.materials.values() {
            let mut t = |x: &Arc<Texture>| add_tex(&mut fs, &mut tex_ids, x);
            let d = match &**m {
                Material::Lambert(l) => SptMaterialDesc { tag: SPT_MAT_LAMBERT, tex: [t(l.albedo()), -1, -1, -1], ior: 1.0, rough_chan: 0, metal_chan: 0 },
                Material::PseudoMaterial(_) => SptMaterialDesc { tag: SPT_MAT_PSEUDO, tex: [-1; 4], ior: 1.0, rough_chan: 0, metal_chan: 0 },
                Material::Conductor(c) => SptMaterialDesc { tag: SPT_MAT_CONDUCTOR, tex: [t(c.ior()), t(c.ior_k()), t(c.roughness_x()), t(c.roughness_y())], ior: 1.0, rough_chan: 0, metal_chan: 0 },
                Material::Dielectric(c) => SptMaterialDesc { tag: SPT_MAT_DIELECTRIC, tex: [t(c.reflectance()), t(c.transmittance()), t(c.roughness_x()), t(c.roughness_y())], ior: c.ior(), rough_chan: 0, metal_chan: 0 },
                Material::Plastic(c) => SptMaterialDesc { tag: SPT_MAT_PLASTIC, tex: [t(c.albedo()), -1, t(c.roughness_x()), t(c.roughness_y())], ior: c.ior(), rough_chan: 0, metal_chan: 0 },
                Material::PbrMetallic(c) => SptMaterialDesc { tag: SPT_MAT_PBR_METALLIC, tex: [t(c.base_color()), t(c.metallic()), t(c.roughness_x()), t(c.roughness_y())], ior: 1.0, rough_chan: c.roughness_chan() as i32, metal_chan: c.metallic_chan() as i32 },
                Material::PbrSpecular(c) => SptMaterialDesc { tag: SPT_MAT_PBR_SPECULAR, tex: [t(c.diffuse()), t(c.specular()), t(c.roughness_x()), t(c.roughness_y())], ior: 1.0, rough_chan: c.roughness_chan() as i32, metal_chan: 0 },
                Material::Subsurface(c) => SptMaterialDesc { tag: SPT_MAT_SUBSURFACE, tex: [t(c.albedo()), t(c.ld()), t(c.roughness_x()), t(c.roughness_y())], ior: c.ior(), rough_chan: 0, metal_chan: 0 },
                _ => panic!("P-NDF materials are out of scope of the wavefront renderer"),
            };
            fs.materials.push(d);
            mat_ids.put(m, fs.materials.len() as i32 - 1);
        }

        // mediums (sigma_s as LOADED, i.e. after Q13), primitives, surfaces: same pattern — push the descriptor, remember the
        // Arc -> index; TriMesh hands over its vertex array and [usize;3] indices in file order (triangle.rs:42-55).
        // ... (one block per component, as in host/scene_loader.cpp)

        // instances in the aggregate's order; xform = trans.trans as 12 column-major floats (instance.rs:19-29)
        for inst in self.instances.values() {
            let a: glam::Affine3A = inst.trans().trans;
            let m = a.matrix3.to_cols_array();
            let t = a.translation;
            let (kind, index) = prim_ids[&Arc::as_ptr(inst.primitive())];
            fs.instances.push(SptInstanceDesc { prim_kind: kind, prim_index: index, surface: /* surface index */ 0,
                                                xform: [m[0], m[1], m[2], m[3], m[4], m[5], m[6], m[7], m[8], t.x, t.y, t.z] });
        }

        // lights in the order of build_light_sampler's Vec (scene_resources.rs:105-138): the `lights` map values
        // (incl. "$env"), then one SPT_LIGHT_SHAPE{instance} per emissive instance; env_light_index as stored (Q8)
        for l in self.lights.values() {
            let _ = match &**l { Light::EnvLight(_) => SPT_LIGHT_ENV, _ => 0 };
            // ... DirLight / PointLight / SpotLight fields -> SptLightDesc
        }
        fs.env_light_index = self.environment_light_index.map_or(-1, |i| i as i32);
        fs.aggregate = if aggregate_type == Some("group") { SPT_AGG_GROUP } else { SPT_AGG_BVH };
        fs.light_sampler = if light_sampler_type == Some("power_is") { SPT_LS_POWER_IS } else { SPT_LS_UNIFORM };
        fs.seal();
        fs
    }
}

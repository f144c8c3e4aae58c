// Flattener: walks a reference-style Scene through its public members only
//   parallel_bvh_node::{left,right,box}            parallel_bvh.h:19-21
//   hitable_list::{list,list_size}                  hitable_list.h:23-24
//   triangle::{V,edge1,edge2,mesh,mat_ptr,inv_area} triangle.h:177-184
//   sphere::{center,radius,mat_ptr}                 sphere.h:20-23
//   material subclasses via dynamic_cast            material.h
//   camera / environment_map                        camera.h:42-48, material.h:206-244
// and produces the POD arrays of include/frt.h: the binary SAH tree collapsed into 4-wide
// nodes with SoA child boxes, triangles in leaf order as {v0,e1,e2} float4 triples, material
// and light tables.  It compiles against this directory's scene_api.h and, unchanged, against
// the reference's own headers (define FRT_WITH_REFERENCE_HEADERS).
#pragma once
#include <cmath>
#include <cstring>
#include <limits>
#include <stdexcept>
#include <unordered_map>
#include <vector>
#include "../../include/frt.h"
#ifndef FRT_WITH_REFERENCE_HEADERS
#include "scene_api.h"
#include "parallel_bvh.h"
#endif

struct frt_flat_scene
{
    std::vector<frt_bvh4_node> nodes;
    std::vector<frt_float4> tris, tri_normals;
    std::vector<frt_sphere> spheres;
    std::vector<frt_material> materials;
    std::vector<frt_light> lights;
    frt_camera camera{};
    float env[3] = {0, 0, 0};
    const float *env_image = nullptr; // HDR environment image, owned by the Scene's image_texture (material.h:209-213)
    uint32_t env_nx = 0, env_ny = 0;
    float ao_t_max = 0.f; // ao.cpp:19-22: 0.5 * size.y of the world's bounding box
    bool any_smooth = false;
    int max_depth_wide = 0;
    // surface textures (texture.h:30-112): table, texels of every image_texture back to back, per-triangle mesh uv
    std::vector<frt_texture> textures;
    std::vector<float> tex_texels, tri_uv;

    frt_scene_desc desc() const
    {
        frt_scene_desc d;
        std::memset(&d, 0, sizeof d);
        d.struct_size = sizeof d;
        d.n_nodes = (uint32_t)nodes.size(); d.nodes = nodes.data();
        d.n_tris = (uint32_t)(tris.size() / 3); d.tris = tris.data();
        d.tri_normals = any_smooth ? tri_normals.data() : nullptr;
        d.n_spheres = (uint32_t)spheres.size(); d.spheres = spheres.data();
        d.n_materials = (uint32_t)materials.size(); d.materials = materials.data();
        d.n_lights = (uint32_t)lights.size(); d.lights = lights.data();
        d.camera = camera;
        std::memcpy(d.env_color, env, sizeof env);
        d.ao_t_max = ao_t_max;
        d.env_image = env_image; d.env_nx = env_image ? env_nx : 0; d.env_ny = env_image ? env_ny : 0;
        d.n_textures = (uint32_t)textures.size(); d.textures = textures.empty() ? nullptr : textures.data();
        d.n_texels = tex_texels.size() / 3; d.tex_texels = tex_texels.empty() ? nullptr : tex_texels.data();
        d.tri_uv = (!textures.empty() && tri_uv.size() == tris.size() * 2) ? tri_uv.data() : nullptr;
        return d;
    }
};

class frt_flattener
{
public:
    // prim_ids: creation-order id per primitive (what the parity tests compare); when empty the
    // ids follow the depth-first leaf order of the tree.  max_leaf: triangles per leaf (1..16); 0 = by scene size —
    // 2 for the reference's small scenes, 4 from 4096 primitives on (measured on B200, profiles/r02_wf_sweep_leaf.log:
    // Cornell 1015 / 1195 / 1044 / 975 Msamples/s at 1 / 2 / 4 / 8 per leaf, the 15 704-triangle mesh scene
    // 985 / 1071 / 1158 / 1091).
    frt_flattener(const std::unordered_map<const hitable *, int> *prim_ids = nullptr, int max_leaf = 0)
        : ids_(prim_ids), asked_leaf_(max_leaf < 1 ? 0 : (max_leaf > 16 ? 16 : max_leaf)), max_leaf_(asked_leaf_) {}

    frt_flat_scene flatten(const Scene *scene)
    {
        out_ = frt_flat_scene();
        bn_.clear(); mat_index_.clear(); tex_index_.clear(); next_id_ = 0;
        pad_ = 0.0;
        int root = scene->world ? import(scene->world.get()) : -1;
        if (root >= 0)
        {
            // conservative float boxes: a few ulps of the largest coordinate the rays can see
            double m = 0;
            for (int a = 0; a < 3; ++a)
                m = std::fmax(m, std::fmax(std::fabs(bn_[root].box.min[a]), std::fabs(bn_[root].box.max[a])));
            for (int a = 0; a < 3; ++a) m = std::fmax(m, std::fabs(scene->cam.origin[a]));
            pad_ = 4e-7 * m;
            max_leaf_ = asked_leaf_ ? asked_leaf_ : (bn_[root].count >= 4096 ? 4 : 2);
            if (bn_[root].prim) { int r2 = (int)bn_.size(); bnode top; top.box = bn_[root].box; top.left = root; top.count = bn_[root].count; top.all_tris = bn_[root].all_tris; bn_.push_back(top); root = r2; }
            emit_wide(root, 1);
        }
        // Scene::lights
        for (int i = 0; i < scene->lights.list_size; ++i) out_.lights.push_back(make_light(scene->lights.list[i]));
        const camera &c = scene->cam;
        put3(out_.camera.origin, c.origin); put3(out_.camera.lower_left_corner, c.lower_left_corner);
        put3(out_.camera.horizontal, c.horizontal); put3(out_.camera.vertical, c.vertical);
        put3(out_.camera.u, c.u); put3(out_.camera.v, c.v); put3(out_.camera.w, c.w);
        out_.camera.lens_radius = (float)c.lens_radius;
        if (scene->env_map && scene->env_map->env_map_tex)
        {
            texture *et = scene->env_map->env_map_tex.get();
            image_texture *it = dynamic_cast<image_texture *>(et);
            if (it && it->img && it->img->type == formats::STBI_HDR && it->img->dataf && it->img->nn == 3)
            {
                out_.env_image = it->img->dataf; // stays owned by the scene, like every other host array
                out_.env_nx = (uint32_t)it->img->nx; out_.env_ny = (uint32_t)it->img->ny;
            }
            else
                put3(out_.env, constant_colour(et, "environment map"));
        }
        if (out_.materials.empty()) { frt_material m{}; m.type = FRT_MAT_LAMBERTIAN; out_.materials.push_back(m); }
        aabb world_box;
        if (scene->world && scene->world->bounding_box(0.0, 0.0, world_box)) out_.ao_t_max = (float)(world_box.size.y() * 0.50f);
        return std::move(out_);
    }

private:
    struct bnode
    {
        aabb box;
        int left = -1, right = -1;       // indices into bn_
        const hitable *prim = nullptr;   // leaf
        int count = 0;                   // primitives below
        bool all_tris = true;
    };

    static void put3(float *dst, const Vector3f &v) { dst[0] = (float)v[0]; dst[1] = (float)v[1]; dst[2] = (float)v[2]; }

    static Vector3f constant_colour(texture *t, const char *what)
    {
        if (!dynamic_cast<constant_texture *>(t))
            throw std::runtime_error(std::string("frt_flattener: only constant_texture is supported for ") + what +
                                     " (emitters and the environment; surface colours may be checker / image textures)");
        hit_record rec;
        rec.u = 0; rec.v = 0;
        return t->value(rec);
    }

    // ---- binary tree import -------------------------------------------------------------------
    int import(const hitable *h)
    {
        if (const parallel_bvh_node *n = dynamic_cast<const parallel_bvh_node *>(h))
        {
            bnode b;
            b.box = n->box;
            const int l = n->left ? import(n->left.get()) : -1;
            const int r = n->right ? import(n->right.get()) : -1;
            if (l < 0 && r < 0) return -1;
            if (l < 0 || r < 0) return l < 0 ? r : l;
            b.left = l; b.right = r;
            b.count = bn_[l].count + bn_[r].count;
            b.all_tris = bn_[l].all_tris && bn_[r].all_tris;
            bn_.push_back(b);
            return (int)bn_.size() - 1;
        }
        if (const hitable_list *hl = dynamic_cast<const hitable_list *>(h))
        {
            // hitable_list::hit is a linear closest-hit loop (hitable_list.cpp:4-21): any tree over
            // the same primitives returns the same hit, so a median split is built here.
            std::vector<int> kids;
            for (int i = 0; i < hl->list_size; ++i) { const int k = import(hl->list[i]); if (k >= 0) kids.push_back(k); }
            return median_tree(kids, 0, (int)kids.size());
        }
        bnode b;
        if (!h->bounding_box(0.0, 0.0, b.box)) throw std::runtime_error("frt_flattener: primitive without a bounding box");
        if (dynamic_cast<const triangle *>(h)) b.all_tris = true;
        else if (dynamic_cast<const sphere *>(h)) b.all_tris = false;
        else throw std::runtime_error("frt_flattener: unsupported hitable (only triangle, sphere, hitable_list, parallel_bvh_node)");
        b.prim = h;
        b.count = 1;
        bn_.push_back(b);
        return (int)bn_.size() - 1;
    }

    int median_tree(std::vector<int> &k, int lo, int hi)
    {
        if (hi - lo <= 0) return -1;
        if (hi - lo == 1) return k[lo];
        aabb box = bn_[k[lo]].box;
        for (int i = lo + 1; i < hi; ++i) box = surrounding_box(box, bn_[k[i]].box);
        int axis = 0;
        for (int a = 1; a < 3; ++a) if (box.size[a] > box.size[axis]) axis = a;
        const int mid = (lo + hi) / 2;
        std::nth_element(k.begin() + lo, k.begin() + mid, k.begin() + hi, [&](int a, int b) {
            return bn_[a].box.min[axis] + bn_[a].box.max[axis] < bn_[b].box.min[axis] + bn_[b].box.max[axis]; });
        bnode b;
        b.box = box;
        b.left = median_tree(k, lo, mid);
        b.right = median_tree(k, mid, hi);
        b.count = bn_[b.left].count + bn_[b.right].count;
        b.all_tris = bn_[b.left].all_tris && bn_[b.right].all_tris;
        bn_.push_back(b);
        return (int)bn_.size() - 1;
    }

    // ---- wide collapse ------------------------------------------------------------------------
    bool is_leaf_group(int i) const { return bn_[i].prim || (bn_[i].all_tris && bn_[i].count <= max_leaf_); }

    uint32_t emit_wide(int bi, int depth)
    {
        if (depth > out_.max_depth_wide) out_.max_depth_wide = depth;
        const uint32_t self = (uint32_t)out_.nodes.size();
        out_.nodes.emplace_back();
        // pull up to four subtrees: always open the widest-area inner candidate
        int kids[4], nk = 0;
        if (bn_[bi].prim) kids[nk++] = bi;
        else { kids[nk++] = bn_[bi].left; if (bn_[bi].right >= 0) kids[nk++] = bn_[bi].right; }
        while (nk < 4)
        {
            int pick = -1;
            double best = -1;
            for (int i = 0; i < nk; ++i)
                if (!is_leaf_group(kids[i]) && bn_[kids[i]].box.area() > best) { best = bn_[kids[i]].box.area(); pick = i; }
            if (pick < 0) break;
            // open in place so that the children (and therefore the triangle array) keep the
            // depth-first, left-to-right order in which parallel_bvh_node::hit visits the leaves
            const int open = kids[pick];
            for (int i = nk; i > pick + 1; --i) kids[i] = kids[i - 1];
            kids[pick] = bn_[open].left;
            kids[pick + 1] = bn_[open].right;
            ++nk;
        }
        frt_bvh4_node node;
        const float inf = std::numeric_limits<float>::infinity();
        for (int k = 0; k < 4; ++k)
        {
            node.lo_x[k] = node.lo_y[k] = node.lo_z[k] = inf;
            node.hi_x[k] = node.hi_y[k] = node.hi_z[k] = -inf;
            node.child[k] = FRT_CHILD_EMPTY;
            node.reserved[k] = 0;
        }
        for (int k = 0; k < nk; ++k)
        {
            const aabb &b = bn_[kids[k]].box;
            node.lo_x[k] = down(b.min[0]); node.hi_x[k] = up(b.max[0]);
            node.lo_y[k] = down(b.min[1]); node.hi_y[k] = up(b.max[1]);
            node.lo_z[k] = down(b.min[2]); node.hi_z[k] = up(b.max[2]);
            if (is_leaf_group(kids[k])) node.child[k] = emit_leaf(kids[k]);
            else node.child[k] = emit_wide(kids[k], depth + 1);
        }
        out_.nodes[self] = node;
        return self;
    }

    float down(double x) const { return std::nextafter((float)(x - pad_), -std::numeric_limits<float>::infinity()); }
    float up(double x) const { return std::nextafter((float)(x + pad_), std::numeric_limits<float>::infinity()); }

    uint32_t emit_leaf(int bi)
    {
        std::vector<const hitable *> prims;
        collect(bi, prims);
        if (const sphere *s = dynamic_cast<const sphere *>(prims[0]))
        {
            const uint32_t first = (uint32_t)out_.spheres.size();
            frt_sphere fs;
            fs.cx = (float)s->center[0]; fs.cy = (float)s->center[1]; fs.cz = (float)s->center[2]; fs.radius = (float)s->radius;
            fs.material = material_index(s->mat_ptr);
            fs.prim_id = (uint32_t)id_of(s);
            fs.pad0 = fs.pad1 = 0;
            out_.spheres.push_back(fs);
            return FRT_CHILD_LEAF | FRT_CHILD_SPHERE | first;
        }
        const uint32_t first = (uint32_t)(out_.tris.size() / 3);
        for (const hitable *h : prims)
        {
            const triangle *t = static_cast<const triangle *>(h);
            const Vector3f &v0 = t->mesh->vertices[t->V[0]];
            const uint32_t flags = ((uint32_t)t->mesh->nTriangles << 1) | (t->mesh->use_geometry_normals ? FRT_TRI_GEOMETRY_NORMALS : 0u);
            out_.tris.push_back(pack(v0, (uint32_t)id_of(t)));
            out_.tris.push_back(pack(t->mesh->vertices[t->V[1]], material_index(t->mat_ptr)));
            out_.tris.push_back(pack(t->mesh->vertices[t->V[2]], flags));
            for (int k = 0; k < 3; ++k) out_.tri_normals.push_back(pack(t->mesh->normals[t->V[k]], 0));
            for (int k = 0; k < 3; ++k) // triangle.h:104-106 interpolates these
            {
                const Vector2f uv = t->mesh->uv ? t->mesh->uv[t->V[k]] : Vector2f(0, 0);
                out_.tri_uv.push_back((float)uv.x); out_.tri_uv.push_back((float)uv.y);
            }
            if (!t->mesh->use_geometry_normals) out_.any_smooth = true;
        }
        return FRT_CHILD_LEAF | ((uint32_t)(prims.size() - 1) << FRT_LEAF_COUNT_SHIFT) | first;
    }

    void collect(int bi, std::vector<const hitable *> &out) const
    {
        if (bn_[bi].prim) { out.push_back(bn_[bi].prim); return; }
        collect(bn_[bi].left, out);
        collect(bn_[bi].right, out);
    }

    static frt_float4 pack(const Vector3f &v, uint32_t w)
    {
        frt_float4 f;
        f.x = (float)v[0]; f.y = (float)v[1]; f.z = (float)v[2];
        std::memcpy(&f.w, &w, 4);
        return f;
    }

    int id_of(const hitable *h)
    {
        if (ids_ && !ids_->empty())
        {
            auto it = ids_->find(h);
            if (it != ids_->end()) return it->second;
        }
        return next_id_++;
    }

    // ---- textures (texture.h:17-112) ----------------------------------------------------------------
    // Index of `t` in the texture table; checker children and image texels are appended on first use.
    uint32_t texture_index(texture *t, int depth)
    {
        auto it = tex_index_.find(t);
        if (it != tex_index_.end()) return it->second;
        if (!t || depth > 6) throw std::runtime_error("frt_flattener: texture missing or nested deeper than 6 checker levels");
        const uint32_t idx = (uint32_t)out_.textures.size();
        out_.textures.emplace_back();
        tex_index_[t] = idx;
        frt_texture ft;
        std::memset(&ft, 0, sizeof ft);
        if (dynamic_cast<constant_texture *>(t))
        {
            const Vector3f c = constant_colour(t, "texture");
            ft.kind = FRT_TEX_CONSTANT;
            ft.p[0] = (float)c[0]; ft.p[1] = (float)c[1]; ft.p[2] = (float)c[2];
        }
        else if (checker_texture *c = dynamic_cast<checker_texture *>(t))
        {
            ft.kind = FRT_TEX_CHECKER;
            ft.a = texture_index(c->tex0, depth + 1);
            ft.b = texture_index(c->tex1, depth + 1);
            ft.p[0] = (float)c->u_scale; ft.p[1] = (float)c->v_scale;
        }
        else if (image_texture *it2 = dynamic_cast<image_texture *>(t))
        {
            const image *img = it2->img.get();
            if (!img || img->nx <= 0 || img->ny <= 0 || img->nn != 3)
                throw std::runtime_error("frt_flattener: image_texture needs a loaded 3-channel image");
            ft.kind = FRT_TEX_IMAGE;
            ft.a = (uint32_t)(out_.tex_texels.size() / 3); ft.b = (uint32_t)img->nx; ft.c = (uint32_t)img->ny;
            const size_t n = (size_t)img->nx * img->ny * 3;
            if (img->type == formats::STBI_HDR)
            {
                if (!img->dataf) throw std::runtime_error("frt_flattener: HDR image_texture without data");
                out_.tex_texels.insert(out_.tex_texels.end(), img->dataf, img->dataf + n);
            }
            else
            {
                if (!img->data) throw std::runtime_error("frt_flattener: image_texture without data");
                for (size_t i = 0; i < n; ++i) out_.tex_texels.push_back((float)FromSrgb(img->data[i] / 255.0)); // texture.h:78-80
            }
        }
        else
            throw std::runtime_error("frt_flattener: unsupported texture class (constant_texture, checker_texture, image_texture)");
        out_.textures[idx] = ft;
        return idx;
    }

    // A material colour: a constant_texture is stored in the material itself (p[at..at+2]); anything else goes to the
    // texture table and the material carries 1 + its index in p[slot] (include/frt.h FRT_MAT_TEX_A / _B).
    void set_colour(frt_material &fm, int at, int slot, texture *t, const char *what)
    {
        if (dynamic_cast<constant_texture *>(t))
        {
            const Vector3f c = constant_colour(t, what);
            fm.p[at] = (float)c[0]; fm.p[at + 1] = (float)c[1]; fm.p[at + 2] = (float)c[2];
            return;
        }
        const uint32_t ref = 1u + texture_index(t, 0);
        std::memcpy(&fm.p[slot], &ref, 4);
    }

    // ---- materials / lights -------------------------------------------------------------------
    uint32_t material_index(material *m)
    {
        auto it = mat_index_.find(m);
        if (it != mat_index_.end()) return it->second;
        frt_material fm;
        std::memset(&fm, 0, sizeof fm);
        auto set3 = [&](int at, const Vector3f &v) { fm.p[at] = (float)v[0]; fm.p[at + 1] = (float)v[1]; fm.p[at + 2] = (float)v[2]; };
        if (lambertian *l = dynamic_cast<lambertian *>(m)) { fm.type = FRT_MAT_LAMBERTIAN; set_colour(fm, 0, FRT_MAT_TEX_A, l->albedo, "lambertian"); }
        else if (modified_phong *p = dynamic_cast<modified_phong *>(m))
        {
            fm.type = FRT_MAT_PHONG;
            set_colour(fm, 0, FRT_MAT_TEX_A, p->diffuse_reflectance, "modified_phong");
            set_colour(fm, 3, FRT_MAT_TEX_B, p->specular_reflectance, "modified_phong");
            fm.p[6] = (float)p->specular_exponent;
        }
        else if (dielectric *d = dynamic_cast<dielectric *>(m))
        {
            fm.type = FRT_MAT_DIELECTRIC;
            set_colour(fm, 0, FRT_MAT_TEX_A, d->specular_reflectance, "dielectric");
            fm.p[6] = (float)d->ref_idx;
        }
        else if (rough_conductor *rc = dynamic_cast<rough_conductor *>(m))
        {
            fm.type = FRT_MAT_ROUGH_CONDUCTOR;
            set_colour(fm, 0, FRT_MAT_TEX_A, rc->specular_reflectance, "rough_conductor");
            set3(3, rc->eta);
            fm.p[6] = (float)rc->alphaU;
            set3(7, rc->k);
            fm.p[10] = rc->distribution_type == microfacet_distributions::ggx ? 0.f : 1.f;
        }
        else if (metal *mt = dynamic_cast<metal *>(m)) { fm.type = FRT_MAT_METAL; set3(0, mt->albedo); }
        else if (diffuse_light *dl = dynamic_cast<diffuse_light *>(m)) { fm.type = FRT_MAT_DIFFUSE_LIGHT; set3(0, constant_colour(dl->emit, "diffuse_light")); }
        else throw std::runtime_error("frt_flattener: unsupported material class");
        const uint32_t idx = (uint32_t)out_.materials.size();
        out_.materials.push_back(fm);
        mat_index_[m] = idx;
        return idx;
    }

    frt_light make_light(const hitable *h)
    {
        frt_light L;
        std::memset(&L, 0, sizeof L);
        auto set = [](float *dst, const Vector3f &v, double w) { dst[0] = (float)v[0]; dst[1] = (float)v[1]; dst[2] = (float)v[2]; dst[3] = (float)w; };
        if (const triangle *t = dynamic_cast<const triangle *>(h))
        {
            L.kind = FRT_LIGHT_TRIANGLE;
            L.material = material_index(t->mat_ptr);
            L.flags = t->mesh->use_geometry_normals ? FRT_TRI_GEOMETRY_NORMALS : 0u;
            const Vector3f &v0 = t->mesh->vertices[t->V[0]];
            set(L.v0, v0, t->inv_area);
            set(L.v1, t->mesh->vertices[t->V[1]], 0);
            set(L.v2, t->mesh->vertices[t->V[2]], 0);
            set(L.n0, t->mesh->normals[t->V[0]], 0);
            set(L.n1, t->mesh->normals[t->V[1]], 0);
            set(L.n2, t->mesh->normals[t->V[2]], 0);
        }
        else if (const sphere *s = dynamic_cast<const sphere *>(h))
        {
            L.kind = FRT_LIGHT_SPHERE;
            L.material = material_index(s->mat_ptr);
            set(L.v0, s->center, s->radius);
        }
        else throw std::runtime_error("frt_flattener: unsupported light primitive");
        return L;
    }

    const std::unordered_map<const hitable *, int> *ids_;
    int asked_leaf_, max_leaf_;
    int next_id_ = 0;
    double pad_ = 0;
    std::vector<bnode> bn_;
    std::unordered_map<material *, uint32_t> mat_index_;
    std::unordered_map<texture *, uint32_t> tex_index_;
    frt_flat_scene out_;
};

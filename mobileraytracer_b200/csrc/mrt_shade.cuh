// Camera rays, sampler draws and the shading functions of the deterministic shaders, in the reference's operation order.
#pragma once
#include "mrt_traverse.cuh"

namespace mrt {

// Perspective::generateRay (Components/Cameras/Perspective.cpp:16-28) / Orthographic::generateRay (Orthographic.cpp:15-23)
MRT_D void generate_ray(const CameraRec &c, float u, float v, float du, float dv, v3 &o, v3 &d) {
    if (c.kind == 0) {
        const float tanValueRight = c.a * (u - 0.5f);
        const float rightFactor = fast_arctan(tanValueRight) + du;
        const v3 right = c.right * rightFactor;
        const float tanValueUp = c.b * (0.5f - v);
        const float upFactor = fast_arctan(tanValueUp) + dv;
        const v3 up = c.up * upFactor;
        const v3 dest = ((c.position + c.direction) + right) + up;
        d = normalize(dest - c.position);
        o = c.position;
    } else {
        const float rightFactor = (u - 0.5f) * c.a;
        const v3 right = c.right * rightFactor + c.right * du;
        const float upFactor = (0.5f - v) * c.b;
        const v3 up = c.up * upFactor + c.up * dv;
        d = c.direction;
        o = (c.position + right) + up;
    }
}

// Renderer::renderScene's tile walk (Renderer.cpp:119-141) in tile space. tile_origin[k] = (startX, startY) of the k-th
// tile of the shuffled tile table, computed on the host with the reference's own integer formula (mrt_api.cu), so odd
// image sizes (where the reference's tiles wrap and overlap) map to the same pixels. A warp covers an 8x4 patch of one
// tile. q = how many pixel-samples a single-threaded reference has rendered before this one: its pixel sampler has
// handed out 2q draws by then (Renderer.cpp:137-138).
MRT_D int tile_patches_x(const RenderParams &rp) { return (rp.bsx + 7) >> 3; }
MRT_D int tile_patches_y(const RenderParams &rp) { return (rp.bsy + 3) >> 2; }
MRT_D bool tile_pixel(const RenderParams &rp, const int2 *__restrict__ tile_origin, int patch, int lane, int sample,
                      int &x, int &y, int &pix, uint32_t &q, int *tile_out = nullptr, int *rank_out = nullptr) {
    const int ptx = tile_patches_x(rp), per_tile = ptx * tile_patches_y(rp);
    const int ksel = patch / per_tile;                        // position among the tiles this handle renders
    if (ksel >= rp.n_tiles) { return false; }
    const int k = rp.tile_begin + ksel * rp.tile_stride;      // position in the reference's shuffled tile walk
    const int r = patch - ksel * per_tile;
    const int lx = (r % ptx) * 8 + (lane & 7), ly = (r / ptx) * 4 + (lane >> 3);
    if (lx >= rp.bsx || ly >= rp.bsy) { return false; }
    const int2 org = __ldg(tile_origin + k);
    x = org.x + lx; y = org.y + ly;
    pix = y * rp.width + x;                                   // Renderer.cpp:149: x may run past the row end, as in the reference
    if (pix >= rp.width * rp.height) { return false; }        // the reference writes out of bounds here; this library does not
    q = (static_cast<uint32_t>(sample) * kNumberOfTiles + static_cast<uint32_t>(k)) * static_cast<uint32_t>(rp.bsx * rp.bsy) +
        static_cast<uint32_t>(ly * rp.bsx + lx);
    if (tile_out) { *tile_out = ksel; }
    if (rank_out) { *rank_out = ly * rp.bsx + lx; }   // valid pixels of a tile are a prefix of its row-major order (pix is monotone in it)
    return true;
}

// The two jitter draws of pixel-sample q (Renderer.cpp:137-138): the pixel sampler's draws 2q and 2q + 1 of the frame, for every
// sampler the reference has (Components/Samplers/*.cpp). The table samplers hand out table[counter++ & mask]
// (Sampler.hpp:58-63); HaltonSeq / Stratified compute the value from the counter and return 1.0 once the counter has
// reached their domain size (HaltonSeq.cpp:10-19, Stratified.cpp:10-19: called with sample = 0, so that is after
// domainSize_ draws); the generators with hidden state (MersenneTwister.cpp:5-11, PCG.cpp:6-13) are served from a stream of
// draws the host took from the sampler object itself (mrt_set_pixel_stream).
MRT_D float pixel_draw(const SceneDev &s, const RenderParams &rp, uint32_t idx) {
    switch (rp.pixel_sampler) {
        case 2: case 3: case 4: return __ldg(s.pixel_table + (idx & kArrayMask));                        // Static{HaltonSeq, MersenneTwister, PCG}
        case 5: return idx < rp.sampler_domain ? halton_sequence(idx, 2u) : 1.0f;                        // HaltonSeq
        case 6: return idx < rp.sampler_domain ? static_cast<float>(idx) / static_cast<float>(rp.sampler_domain) : 1.0f;   // Stratified
        case 7: return __ldg(s.pixel_table + (idx < rp.pixel_stream_n ? idx : idx % rp.pixel_stream_n)); // caller's stream
        default: return 0.5f;                                                                            // Constant(0.5)
    }
}

MRT_D uint32_t hash_u32(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t h = a * 0x9E3779B1u ^ (b + 0x7F4A7C15u) * 0x85EBCA77u ^ (c + 0x165667B1u) * 0xC2B2AE3Du;
    h ^= h >> 16; h *= 0x85EBCA6Bu; h ^= h >> 13; h *= 0xC2B2AE35u; h ^= h >> 16;
    return h;
}

// Where a draw of the shared tables comes from. The reference hands out table[counter++ & mask] in execution order
// (Shader.cpp:188-233, Sampler.hpp:58-63). In replay mode the exact counters are supplied; otherwise every
// (pixel-sample q, event e, stream) gets its own decorrelated slot of the same table.
struct DrawState {
    uint32_t q;         // pixel-sample sequence id
    uint32_t e;         // running event counter of this pixel-sample
    uint32_t seed;
    int32_t mode;       // ReplayMode
    uint32_t a_next;    // value of Shader::getLightIndex's counter (Shader.cpp:224) at this pixel-sample's next draw
    uint32_t l_next[kReplayMaxLights];   // per light: value of its sampler's counter (AreaLight.cpp:18-19) at the next draw
};
enum DrawStream : uint32_t { DS_LIGHT_INDEX = 1, DS_LIGHT_POS = 2, DS_HEMI = 3, DS_RR = 4 };

MRT_D float draw(const float *table, DrawState &ds, uint32_t stream) {
    const uint32_t i = hash_u32(ds.q, ds.e++, stream ^ ds.seed) & kArrayMask;
    return __ldg(table + i);
}

struct Surface {
    v3 point, normal;
    v3 Le, Kd, Ks, Kt;
    float ior;
    int32_t texture;             // the material's texture or -1
    int32_t skipKind, skipIdx;   // the primitive a ray leaving this surface must skip (Ray::primitive_)
};

// matp: the material records as float4 quadruples, in global or (wavefront shade kernel) shared memory
MRT_D void load_material(const float4 *matp, int m, Surface &sf) {
    const float4 *mp = matp + 4 * m;
    const float4 a = mp[0], b = mp[1], c = mp[2], d = mp[3];
    sf.Le = mk3(a.x, a.y, a.z); sf.Kd = mk3(a.w, b.x, b.y); sf.Ks = mk3(b.z, b.w, c.x); sf.Kt = mk3(c.y, c.z, c.w); sf.ior = d.x;
    sf.texture = __float_as_int(d.y);
}

// Texture::loadColor (Texture.cpp:37-48): nearest texel, no wrap (the loader already took fract() of the coordinates,
// OBJLoader.cpp:193-212, Utils.cpp:177-180), bytes index..index+2 whatever the channel count. The reference reads past
// the image when a coordinate reaches 1.0; here the index is clamped to the texture's own bytes.
MRT_D v3 texture_load_color(const SceneDev &s, int tex, float tu, float tv) {
    const TextureRec tr = s.textures[tex];
    const int32_t u = static_cast<int32_t>(tu * static_cast<float>(tr.width));
    const int32_t v = static_cast<int32_t>(tv * static_cast<float>(tr.height));
    uint32_t index = static_cast<uint32_t>(v * tr.width * tr.channels + u * tr.channels);
    if (index + 3u > tr.bytes) { index = tr.bytes >= 3u ? tr.bytes - 3u : 0u; }
    const uint8_t *p = s.texels + tr.offset + index;
    return mk3(static_cast<float>(p[0]) / 255.0f, static_cast<float>(p[1]) / 255.0f, static_cast<float>(p[2]) / 255.0f);
}

// What the primitive's intersect() stores in the Intersection (Triangle.cpp:100-108, Sphere.cpp:69-78, Plane.cpp:64-70)
// plus Shader::rayTrace's material lookup (Shader.cpp:112-121; AreaLight.cpp:32-41 for lights).
MRT_D void resolve_surface(const SceneDev &s, v3 o, v3 d, const HitRec &h, Surface &sf, const float4 *matp = nullptr, const float4 *lightp = nullptr) {
    if (matp == nullptr) { matp = reinterpret_cast<const float4 *>(s.materials); }
    if (lightp == nullptr) { lightp = reinterpret_cast<const float4 *>(s.lights); }
    sf.point = o + d * h.t;
    sf.skipKind = HIT_NONE; sf.skipIdx = -1;
    if (h.kind == HIT_TRIANGLE) {
        const float4 *sp = reinterpret_cast<const float4 *>(s.tri_shade + h.idx);
        const float4 s0 = __ldg(sp), s1 = __ldg(sp + 1), s2 = __ldg(sp + 2), s3 = __ldg(sp + 3);
        const float w = 1.0f - h.u - h.v;
        const v3 nA = mk3(s0.x, s0.y, s0.z), nB = mk3(s0.w, s1.x, s1.y), nC = mk3(s1.z, s1.w, s2.x);
        sf.normal = normalize((nA * w + nB * h.u) + nC * h.v);
        load_material(matp, __float_as_int(s3.w), sf);
        if (sf.texture >= 0) {
            // Triangle.cpp:98 texCoords = tA * w + tB * u + tC * v; Shader.cpp:116-120: both >= 0 -> Kd comes from the texture
            const float tu = (s2.y * w + s2.w * h.u) + s3.y * h.v, tv = (s2.z * w + s3.x * h.u) + s3.z * h.v;
            if (tu >= 0.0f && tv >= 0.0f) { sf.Kd = texture_load_color(s, sf.texture, tu, tv); }
        }
        sf.skipKind = HIT_TRIANGLE; sf.skipIdx = h.idx;
    } else if (h.kind == HIT_SPHERE) {
        const SphereRec sp = s.spheres[h.idx];
        sf.normal = normalize(sf.point - mk3(sp.cx, sp.cy, sp.cz));
        load_material(matp, sp.material, sf);
    } else if (h.kind == HIT_PLANE) {
        const PlaneRec p = s.planes[h.idx];
        sf.normal = mk3(p.nx, p.ny, p.nz);
        load_material(matp, p.material, sf);
        sf.skipKind = HIT_PLANE; sf.skipIdx = h.idx;
    } else {  // HIT_LIGHT: material = the light's radiance_ (Le only)
        // LightRec as four float4: kind A.xyz | AB.xyz AC.x | AC.yz Le.xy | Le.z - - -
        const float4 l1 = lightp[4 * h.idx + 1], l2 = lightp[4 * h.idx + 2], l3 = lightp[4 * h.idx + 3];
        // the light triangle's face normal (Triangle::Builder, Triangle.cpp:328-340); no shader reads it: Le > 0 ends the path
        sf.normal = normalize(normalize(cross(mk3(l1.w, l2.x, l2.y), mk3(l1.x, l1.y, l1.z))));
        sf.Le = mk3(l2.z, l2.w, l3.x);
        sf.Kd = sf.Ks = sf.Kt = mk3(0.0f, 0.0f, 0.0f);
        sf.ior = 1.0f;
    }
}

// Shader::getLightIndex (Shader.cpp:223-233)
MRT_D uint32_t light_index(const SceneDev &s, DrawState &ds) {
    const float r = ds.mode == REPLAY_OFF ? draw(s.shader_table, ds, DS_LIGHT_INDEX)
                                          : __ldg(s.shader_table + ((ds.a_next++) & kArrayMask));
    return static_cast<uint32_t>(floorf(r * static_cast<float>(static_cast<uint32_t>(s.n_lights)) * 0.99999f));
}
// PointLight::getPosition (PointLight.cpp:16-18) / AreaLight::getPosition (AreaLight.cpp:17-26)
MRT_D v3 light_position(const SceneDev &s, const LightRec &l, DrawState &ds, uint32_t chosen) {
    if (l.kind == 0) { return mk3(l.A[0], l.A[1], l.A[2]); }
    float r, q;
    if (ds.mode == REPLAY_OFF) {
        r = draw(s.static_halton, ds, DS_LIGHT_POS);
        q = draw(s.static_halton, ds, DS_LIGHT_POS);
    } else {
        uint32_t c = 0u;   // l_next[chosen] without dynamic indexing (keeps the array in registers)
#pragma unroll
        for (int k = 0; k < kReplayMaxLights; ++k) { if (static_cast<uint32_t>(k) == chosen) { c = ds.l_next[k]; ds.l_next[k] = c + 2u; } }
        r = __ldg(s.static_halton + (c & kArrayMask));
        q = __ldg(s.static_halton + ((c + 1u) & kArrayMask));
    }
    if (r + q >= 1.0f) { r = 1.0f - r; q = 1.0f - q; }
    return (mk3(l.A[0], l.A[1], l.A[2]) + r * mk3(l.AB[0], l.AB[1], l.AB[2])) + q * mk3(l.AC[0], l.AC[1], l.AC[2]);
}

struct RayCounters { uint32_t primary, shadow, secondary; };

// Direct lighting term shared by Whitted (Whitted.cpp:37-64) and NoShadows (NoShadows.cpp:25-43); SHADOWS selects which.
template <bool SHADOWS>
MRT_D v3 direct_light(const SceneDev &s, const RenderParams &rp, const Surface &sf, DrawState &ds, RayCounters &rc, int *overflow) {
    v3 rgb = mk3(0.0f, 0.0f, 0.0f);
    if (has_positive(sf.Kd) && s.n_lights > 0) {
        if (ds.mode == REPLAY_COUNT_INDEX) { ds.a_next += static_cast<uint32_t>(rp.spl); return rgb; }
        for (int i = 0; i < rp.spl; ++i) {
            const uint32_t chosen = light_index(s, ds);
            const LightRec &l = s.lights[chosen];
            const v3 lightPosition = light_position(s, l, ds, chosen);
            if (ds.mode == REPLAY_COUNT_LIGHTS) { continue; }
            v3 vectorToLight = lightPosition - sf.point;
            const float distanceToLight = length(vectorToLight);
            vectorToLight = normalize(vectorToLight);
            const float cosNl = dot(sf.normal, vectorToLight);
            if (cosNl > 0.0f) {
                bool lit = true;
                if (SHADOWS) {
                    ++rc.shadow;
                    lit = !trace_shadow(s, sf.point, vectorToLight, distanceToLight, sf.skipKind, sf.skipIdx, overflow);
                }
                if (lit) { rgb = rgb + mk3(l.Le[0], l.Le[1], l.Le[2]) * cosNl; }
            }
        }
        rgb = rgb * sf.Kd;
        rgb = rgb / static_cast<float>(rp.spl);
    }
    return rgb;
}

// DepthMap::shade (DepthMap.cpp:12-17)
MRT_D v3 shade_depthmap(const RenderParams &rp, v3 rayOrigin, float t) {
    const float maxDist = length(rp.depth_max_point - rayOrigin) * 1.1f;
    const float depth = std_max((maxDist - t) / maxDist, 0.0f);
    return mk3(depth, depth, depth);
}
// DiffuseMaterial::shade (DiffuseMaterial.cpp:11-27)
MRT_D v3 shade_diffuse(const Surface &sf) {
    if (has_positive(sf.Kd)) { return sf.Kd; }
    if (has_positive(sf.Ks)) { return sf.Ks; }
    if (has_positive(sf.Kt)) { return sf.Kt; }
    if (has_positive(sf.Le)) { return sf.Le; }
    return mk3(0.0f, 0.0f, 0.0f);
}

// Whitted::shade driven by Shader::rayTrace (Whitted.cpp:13-93, Shader.cpp:86-123): the recursion is unrolled into an
// explicit frame stack so that every partial sum is formed in the reference's order.
struct WhittedFrame {
    v3 rgb, Kd, Ks, Kt, P, N, D;
    float ior;
    int32_t skipKind, skipIdx, depth, stage;
};

template <int SHADER>
MRT_D v3 ray_trace(const SceneDev &s, const RenderParams &rp, v3 o, v3 d, int depth0, DrawState &ds, RayCounters &rc, int *overflow) {
    HitRec h;
    Surface sf;
    if (SHADER != 1) {
        trace_closest(s, o, d, HIT_NONE, -1, h, overflow);
        if (h.kind == HIT_NONE) { return mk3(0.0f, 0.0f, 0.0f); }
        if (SHADER == 3) { return shade_depthmap(rp, o, h.t); }
        resolve_surface(s, o, d, h, sf);
        if (SHADER == 4) { return shade_diffuse(sf); }
        // NoShadows (NoShadows.cpp:12-46)
        if (has_positive(sf.Le)) { return sf.Le; }
        v3 rgb = direct_light<false>(s, rp, sf, ds, rc, overflow);
        return rgb + sf.Kd * 0.1f;
    }
    WhittedFrame fr[kRayDepthMax + 1];
    int sp = 0;
    int depth = depth0;
    int skipKind = HIT_NONE, skipIdx = -1;
    v3 result = mk3(0.0f, 0.0f, 0.0f);
    bool calling = true;
    while (true) {
        if (calling) {
            result = mk3(0.0f, 0.0f, 0.0f);
            calling = false;
            if (depth > kRayDepthMax) { continue; }   // Whitted.cpp:15-17: rgb stays 0 whatever the ray hits
            trace_closest(s, o, d, skipKind, skipIdx, h, overflow);
            if (h.kind == HIT_NONE) { continue; }
            resolve_surface(s, o, d, h, sf);
            if (has_positive(sf.Le)) { result = sf.Le; continue; }
            v3 rgb = direct_light<true>(s, rp, sf, ds, rc, overflow);
            const bool spec = has_positive(sf.Ks), trans = has_positive(sf.Kt);
            if (!spec && !trans) { result = rgb + sf.Kd * 0.1f; continue; }
            WhittedFrame &f = fr[sp++];
            f.rgb = rgb; f.Kd = sf.Kd; f.Ks = sf.Ks; f.Kt = sf.Kt; f.P = sf.point; f.N = sf.normal; f.D = d;
            f.ior = sf.ior; f.skipKind = sf.skipKind; f.skipIdx = sf.skipIdx; f.depth = depth;
            ++rc.secondary;
            if (spec) { f.stage = 0; d = reflect(f.D, f.N); }
            else { f.stage = 1; d = refract(f.D, f.N, 1.0f / f.ior); }
            o = f.P; depth = f.depth + 1; skipKind = f.skipKind; skipIdx = f.skipIdx;
            calling = true;
        } else {
            if (sp == 0) { return result; }
            WhittedFrame &f = fr[sp - 1];
            if (f.stage == 0) {
                f.rgb = f.rgb + f.Ks * result;
                if (has_positive(f.Kt)) {
                    f.stage = 1;
                    ++rc.secondary;
                    d = refract(f.D, f.N, 1.0f / f.ior);
                    o = f.P; depth = f.depth + 1; skipKind = f.skipKind; skipIdx = f.skipIdx;
                    calling = true;
                    continue;
                }
            } else {
                f.rgb = f.rgb + f.Kt * result;
            }
            result = f.rgb + f.Kd * 0.1f;
            --sp;
        }
    }
}

}  // namespace mrt

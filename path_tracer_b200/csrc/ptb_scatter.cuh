// One path vertex: medium transport, nested-dielectric priority, BSDF / sky-lobe
// sampling, hero-wavelength bookkeeping, sky emission, path regeneration.
//
// Reference behaviour: src/integrator/basic_scatter.glsl (GenerateNewPath :7-42,
// ResolveMedium :44-64, SampleSurfaceIntegrand :68-109, Scatter :114-310,
// main :312-360) with the state quantisation of src/integrator/basic.glsl.inc
// (:87-225: oct-encoded directions, 16-bit shape ids).  Random draws happen in
// the reference's order.
//
// What is different is where the state lives: PathState/Ray/SurfaceHit are plain
// register structs, so the same code serves the wavefront kernels (state in
// 16-byte SoA records in HBM) and the megakernel (state in registers across
// rounds, quantised through the same pack/unpack so both produce equal images).
#pragma once
#include "ptb_bsdf.cuh"
#include "ptb_trace.cuh"

namespace ptb {

// Push constants of a dispatch (basic.glsl.inc:76-85) + film geometry.
struct RenderConsts {
    u32 camera_index, render_flags, path_length_limit;
    float termination_probability;
    u32 random_seed;
    u32 width, height, tiles_x;
    u32 slot_count;         // slots covered by this launch (multiple of 256)
    u32 slot_base;          // first slot of this launch: a film may be rendered as several slot ranges (pipelines)
    u32 width_magic, tiles_x_magic;   // floor(2^32 / width), floor(2^32 / tiles_x): see udivmod (set_film_geometry fills them)
};

// n / d and n % d for a launch-uniform divisor with m = floor(2^32 / d) (0xFFFFFFFF for d = 1): the high half of
// n * m is the quotient or one less, for every 32-bit n, so one correction step makes it exact - a multiply-high and a
// few ALU operations instead of the ~20-instruction general unsigned division.
PTB_HD void udivmod(u32 n, u32 d, u32 m, u32& q, u32& r) {
#ifdef __CUDA_ARCH__
    q = __umulhi(n, m);
#else
    q = (u32)(((unsigned long long)n * m) >> 32);
#endif
    r = n - q * d;
    if (r >= d) { q += 1u; r -= d; }
}
inline void set_film_geometry(RenderConsts& pc, u32 width, u32 height) {
    pc.width = width; pc.height = height;
    pc.tiles_x = (width + 15u) / 16u;
    pc.width_magic = width > 1u ? (u32)(0x100000000ull / width) : 0xFFFFFFFFu;
    pc.tiles_x_magic = pc.tiles_x > 1u ? (u32)(0x100000000ull / pc.tiles_x) : 0xFFFFFFFFu;
}

struct PathState {
    u32 pixel;              // y * width + x
    float lambda0;
    f4 throughput, probability;
    f3 sample;
    u32 active[4];          // nested-dielectric shape stack, PTB_SHAPE_NONE = empty
    u32 length;             // vertices processed (PathLengthLimit generalisation)
};

// What scatter sees of the closest hit after the trace->scatter quantisation.
struct SurfaceHit {
    float t;
    u32 shape, material;    // shape == PTB_SHAPE_NONE: miss
    f3 position, normal, tangent_x, tangent_y;
    f2 uv;
};

// The 24-byte wavefront hit record (basic.glsl.inc:142-157).
struct PackedHit { float t; u32 shape_material, normal, tangent; float u, v; };

PTB_HD PackedHit pack_hit(const HitId& h, const Surface& sf) {
    PackedHit p;
    p.t = h.t; p.shape_material = (h.shape << 16) | sf.material;
    p.normal = pack_unit_vector(sf.normal); p.tangent = pack_unit_vector(sf.tangent_x);
    p.u = sf.uv.x; p.v = sf.uv.y;
    return p;
}
PTB_HD PackedHit packed_miss() { PackedHit p; p.t = PTB_HIT_TIME_LIMIT; p.shape_material = 0xFFFFFFFFu; p.normal = p.tangent = 0; p.u = p.v = 0; return p; }

// basic.glsl.inc:99-131
PTB_HD SurfaceHit unpack_hit(const PackedHit& p, const Ray& ray) {
    SurfaceHit h;
    if (p.shape_material == 0xFFFFFFFFu) {
        h.shape = PTB_SHAPE_NONE; h.material = 0; h.t = PTB_HIT_TIME_LIMIT;
        h.position = h.normal = h.tangent_x = h.tangent_y = mk3(0, 0, 0); h.uv = mk2(0, 0);
        return h;
    }
    h.shape = p.shape_material >> 16;
    h.material = p.shape_material & 0xFFFFu;
    h.t = p.t;
    h.normal = unpack_unit_vector(p.normal);
    h.tangent_x = unpack_unit_vector(p.tangent);
    h.tangent_y = cross(h.normal, h.tangent_x);
    h.uv = mk2(p.u, p.v);
    h.position = ray.origin + h.t * ray.velocity;
    return h;
}

// basic.glsl.inc:183-195,214-215: the active-shape stack as stored (2 x u16|u16).
// The store ORs a full 32-bit NONE into the low half, so an empty entry 0/2 also
// clears entry 1/3 — part of the reference's behaviour, kept.
PTB_HD void pack_active(const u32* a, u32& w01, u32& w23) { w01 = (a[1] << 16) | a[0]; w23 = (a[3] << 16) | a[2]; }
PTB_HD void unpack_active(u32 w01, u32 w23, u32* a) {
    a[0] = w01 & 0xFFFFu; a[1] = w01 >> 16; a[2] = w23 & 0xFFFFu; a[3] = w23 >> 16;
#pragma unroll
    for (int k = 0; k < 4; k++) if (a[k] == 0xFFFFu) a[k] = PTB_SHAPE_NONE;
}

// basic_scatter.glsl:7-42
PTB_HD void generate_new_path(const DeviceScene& s, const RenderConsts& pc, Rng& g, u32 x, u32 y, PathState& path, Ray& ray) {
    f2 sp = mk2((float)(int)x, (float)(int)y);
    if (pc.render_flags & PTB_RENDER_FLAG_SAMPLE_JITTER) {
        float jx = g.uniform();
        float jy = g.uniform();
        sp = sp + mk2(jx, jy);
    } else {
        sp = sp + mk2(0.5f, 0.5f);
    }
    f2 nsp = sp / mk2((float)pc.width, (float)pc.height);
    ray = generate_camera_ray(s, pc.camera_index, nsp, g);
    path.pixel = y * pc.width + x;
    path.lambda0 = g.uniform();
    path.throughput = mk4(1.0f);
    path.probability = mk4(1.0f);
    path.sample = mk3(0.0f);
#pragma unroll
    for (int k = 0; k < 4; k++) path.active[k] = PTB_SHAPE_NONE;
    path.length = 0;
}

// basic_scatter.glsl:44-64.  A shape whose material defines no medium leaves
// the reference's struct uninitialised; pinned to vacuum here (same as the oracle).
PTB_HDN Medium resolve_medium(const DeviceScene& s, u32 shape, f4 lambda) {
    Medium m;
    m.ior = mk4(1.0f); m.absorption = mk4(0.0f); m.scattering = mk4(0.0f); m.anisotropy = 0.0f;
    if (shape == PTB_SHAPE_NONE) {
        m.priority = 0xFFFFFFFFu;
        if (s.features & PTB_FEAT_SCENE_SCATTER) m.scattering = mk4(s.globals.SceneScatterRate);
    } else {
        material_load_medium(s, PTB_LDG(&s.shapes[shape].MaterialIndex), lambda, m);
        m.priority = shape;
    }
    return m;
}

// basic_scatter.glsl:68-109.  Integrator naming: `out` points toward the viewer, `in` is sampled.
PTB_HD bool sample_surface(const DeviceScene& s, Rng& g, const SurfaceHit& hit, const BsdfParams& p, f3 out, f3& in, f4& throughput, f4& pdf) {
    const float light_p = material_has_dirac(s, p) ? 0.0f : s.globals.SkyboxSamplingProbability;
    f4 material_pdf = mk4(0.0f);
    const f3 mean = mk3(s.globals.SkyboxMeanDirection[0], s.globals.SkyboxMeanDirection[1], s.globals.SkyboxMeanDirection[2]);
    const f3 sky_mean = mk3(dot(mean, hit.tangent_x), dot(mean, hit.tangent_y), dot(mean, hit.normal));
    if (g.uniform() < light_p) {
        in = random_vmf(g, s.globals.SkyboxConcentration, sky_mean);
        if (in.z < 0.0f) return false;
        if (!material_eval(s, p, out, in, throughput, material_pdf)) return false;
    } else {
        if (!material_sample(s, p, out, g, in, throughput, material_pdf)) return false;
    }
    const f4 sky_pdf = mk4(vmf_pdf(s.globals.SkyboxConcentration, sky_mean, in));
    pdf = light_p * sky_pdf + (1.0f - light_p) * material_pdf;
    return true;
}

// basic_scatter.glsl:114-310.  Returns true when `ray` holds the extension ray.
PTB_HD bool scatter_vertex(const DeviceScene& s, const RenderConsts& pc, Rng& g, PathState& path, Ray& ray, const SurfaceHit& hit) {
    const float l0 = path.lambda0;
    const f4 lambda = mk4(gmix(PTB_LAMBDA_MIN, PTB_LAMBDA_MAX, l0),
                          gmix(PTB_LAMBDA_MIN, PTB_LAMBDA_MAX, gfract(l0 + 0.25f)),
                          gmix(PTB_LAMBDA_MIN, PTB_LAMBDA_MAX, gfract(l0 + 0.50f)),
                          gmix(PTB_LAMBDA_MIN, PTB_LAMBDA_MAX, gfract(l0 + 0.75f)));

    u32 active = PTB_SHAPE_NONE;      // highest-priority (lowest index) enclosing shape
#pragma unroll
    for (int k = 0; k < 4; k++) if (path.active[k] < active) active = path.active[k];

    const Medium medium = resolve_medium(s, active, lambda);

    // without media, absorption = scattering = 0: exp(-0 * t) = 1 exactly and no scattering event is drawn
    const bool media = (s.features & (PTB_FEAT_TRANSLUCENT | PTB_FEAT_SCENE_SCATTER)) != 0u;
    if (media) path.throughput = path.throughput * exp4((-medium.absorption) * hit.t);

    float scatter_t = PTB_HIT_TIME_LIMIT;
    if (media && medium.scattering.x > 0.0f) scatter_t = (-logf(g.uniform())) / medium.scattering.x;

    if (hit.t >= scatter_t) {
        if (scatter_t < PTB_HIT_TIME_LIMIT) {
            // volume scattering event
            ray.origin = ray.origin + ray.velocity * scatter_t;
            f3 fx, fy, fz = ray.velocity;
            coordinate_frame(fz, fx, fy);
            const float u1 = g.uniform();
            const float u2 = g.uniform();
            const f3 sc = sample_hg(medium.anisotropy, u1, u2);
            f4 density = medium.scattering * exp4((-medium.scattering) * scatter_t);
            density = density / fmaxf(PTB_EPSILON, hmax4(density));
            path.throughput = path.throughput * density;
            path.probability = path.probability * density;
            ray.velocity = normalize((fx * sc.x + fy * sc.y) + fz * sc.z);
            ray.duration = PTB_HIT_TIME_LIMIT;
        } else {
            // escaped: the sky is the only emitter
            const f4 emission = skybox_radiance(s, ray.velocity, lambda);
            const float cluster_pdf = ((path.probability.x + path.probability.y) + path.probability.z) + path.probability.w;
            path.sample = path.sample + observer_times(lambda, emission * path.throughput) / cluster_pdf;
            path.probability = mk4(0.0f);
        }
        return hmax4(path.probability) > PTB_EPSILON;
    }

    f3 in;
    const f3 out = -mk3(dot(ray.velocity, hit.tangent_x), dot(ray.velocity, hit.tangent_y), dot(ray.velocity, hit.normal));

    bool real_surface;
    f4 exterior_ior = mk4(1.0f);
    if (out.z > 0.0f) {
        real_surface = medium.priority > hit.shape;
        if (real_surface) exterior_ior = medium.ior;
    } else {
        real_surface = medium.priority == hit.shape;
        if (real_surface) {
            u32 exterior = PTB_SHAPE_NONE;
#pragma unroll
            for (int k = 0; k < 4; k++) {
                if (path.active[k] == active) continue;
                if (path.active[k] < exterior) exterior = path.active[k];
            }
            exterior_ior = resolve_medium(s, exterior, lambda).ior;
        }
    }

    if (real_surface) {
        BsdfParams p;
        p.material = hit.material; p.uv = hit.uv; p.lambda = lambda; p.exterior_ior = exterior_ior;
        f4 throughput = mk4(0.0f), pdf = mk4(0.0f);
        if (!sample_surface(s, g, hit, p, out, in, throughput, pdf)) return false;
        const float scale = 1.0f / fmaxf(PTB_EPSILON, hmax4(pdf));
        path.throughput = path.throughput * (throughput * scale);
        path.probability = path.probability * (pdf * scale);
    } else {
        in = -out;
    }

    if (in.z * out.z < 0.0f) {
        // crossed the boundary: push on entry, pop on exit
        if (out.z > 0.0f) {
            bool done = false;
#pragma unroll
            for (int k = 0; k < 4; k++) if (!done && path.active[k] == PTB_SHAPE_NONE) { path.active[k] = hit.shape; done = true; }
        } else {
            bool done = false;
#pragma unroll
            for (int k = 0; k < 4; k++) if (!done && path.active[k] == hit.shape) { path.active[k] = PTB_SHAPE_NONE; done = true; }
        }
    }

    if (g.uniform() < pc.termination_probability) return false;
    path.probability = path.probability * (1.0f - pc.termination_probability);

    ray.velocity = (in.x * hit.tangent_x + in.y * hit.tangent_y) + in.z * hit.normal;
    ray.origin = hit.position + 1e-3f * ray.velocity;
    ray.duration = PTB_HIT_TIME_LIMIT;
    return hmax4(path.probability) > PTB_EPSILON;
}

// One scatter invocation for the pixel (x, y) that owns this slot
// (basic_scatter.glsl:312-360, non-restart branch).  `accum` is the film's
// RGBA32F pixel (rgb = sum of XYZ samples, a = sample count); exactly one slot
// owns each pixel, so the read-modify-write needs no atomics.  Returns true if a
// path finished (for the Mpaths/s counter).
PTB_HD bool scatter_invocation(const DeviceScene& s, const RenderConsts& pc, u32 x, u32 y, PathState& path, Ray& ray,
                               const SurfaceHit& hit, float4* accum) {
    Rng g;
    g.state = y * 65537u + x + pc.random_seed * 277803737u;
    bool alive = scatter_vertex(s, pc, g, path, ray, hit);
    path.length += 1;
    if (alive && pc.path_length_limit != 0 && path.length >= pc.path_length_limit) alive = false;
    if (alive) return false;
    float4* px = accum + path.pixel;
    float4 v = make_float4(path.sample.x, path.sample.y, path.sample.z, 1.0f);
    if (pc.render_flags & PTB_RENDER_FLAG_ACCUMULATE) {
        float4 old = *px;
        v.x += old.x; v.y += old.y; v.z += old.z; v.w += old.w;
    }
    *px = v;
    u32 px_x, px_y;
    udivmod(path.pixel, pc.width, pc.width_magic, px_y, px_x);
    generate_new_path(s, pc, g, px_x, px_y, path, ray);
    return true;
}

}  // namespace ptb

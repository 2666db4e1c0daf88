/* TEST INFRASTRUCTURE — CPU oracle of the agent-view ray caster.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this; the product never does.
 *
 * The reference produces these images either with a Vulkan rasteriser (unavailable on this box)
 * or with Madrona's CUDA software ray caster, which is device code inside the NVRTC megakernel
 * and cannot be built or run here.  What CAN be run here is the reference's host-compilable mesh
 * ray tracing, MeshBVH::traceRay (M/include/madrona/mesh_bvh.inl:79-168, 319-378, 380-462): when
 * the test installs it with hs_oracle_set_tracer (tests/views_util.py takes it from
 * oracle/_ref/libhs_ref.so, hsref_mesh_trace), the per-instance ray / triangle stage below IS
 * reference code and only these parts remain a restatement (PARITY UNPINNED for them):
 *   camera ray generation      calculateOutRay, M/src/mw/device/bvh_raycast.cpp:58-88
 *   the instance loop          ray -> object space, t rescale, closest instance, bvh_raycast.cpp:614-633
 *   shading                    0.2 x material colour x texture, bvh_raycast.cpp:783-797, 822-935
 * Without a tracer installed the ray / triangle stage falls back to the brute-force restatement
 * further down (every triangle of every instance, no acceleration structure), following
 *   calculateOutRay            M/src/mw/device/bvh_raycast.cpp:58-88
 *   computeRayIsectInfo        bvh_raycast.cpp:228-270
 *   rayTriangleIntersection    bvh_raycast.cpp:317-447 (MADRONA_MESHBVH_BACKFACE_CULLING on,
 *                              M/include/madrona/mesh_bvh.inl:3)
 *   instance transform         bvh_raycast.cpp:614-633
 *   computeFragment / writes   bvh_raycast.cpp:822-935, 1003-1020 (no lights in this sim)
 *   camera                     src/sim.cpp:1399-1403, M/src/render/ecs_system.cpp:274-313, 655-660
 * Build: oracle/Makefile -> oracle/_ref/libviews_oracle.so (gcc -O2 -ffp-contract=off). */
#include <math.h>
#include <stdint.h>
#include <string.h>

typedef struct { float x, y, z; } v3;
typedef struct { float w, x, y, z; } quat;

static v3 add(v3 a, v3 b) { v3 r = {a.x + b.x, a.y + b.y, a.z + b.z}; return r; }
static v3 sub(v3 a, v3 b) { v3 r = {a.x - b.x, a.y - b.y, a.z - b.z}; return r; }
static v3 scl(v3 a, float s) { v3 r = {a.x * s, a.y * s, a.z * s}; return r; }
static v3 crs(v3 a, v3 b) { v3 r = {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; return r; }
static float len(v3 a) { return sqrtf(a.x * a.x + a.y * a.y + a.z * a.z); }
static v3 nrm(v3 a) { return scl(a, 1.f / len(a)); }
static float cmp(v3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }
/* Quat::rotateVec, M/include/madrona/math.inl:717-726 */
static v3 rot(quat q, v3 v)
{
    v3 p = {q.x, q.y, q.z};
    v3 pxv = crs(p, v);
    v3 pxpxv = crs(p, pxv);
    return add(v, scl(add(scl(pxv, q.w), pxpxv), 2.f));
}
static quat qconj(quat q) { quat r = {q.w, -q.x, -q.y, -q.z}; return r; }

typedef struct { int kx, ky, kz; float Sx, Sy, Sz; } isect;

static isect isect_info(v3 d)
{
    float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    int kz;
    if (ax > ay && ax > az) kz = 0; else if (ay > az) kz = 1; else kz = 2;
    int kx = kz + 1; if (kx == 3) kx = 0;
    int ky = kx + 1; if (ky == 3) ky = 0;
    if (cmp(d, kz) < 0.f) { int t = kx; kx = ky; ky = t; }
    float inv = 1.f / cmp(d, kz);
    isect r = {kx, ky, kz, cmp(d, kx) * inv, cmp(d, ky) * inv, inv};
    return r;
}

static int tri_hit(const float *t, isect R, v3 org, float t_max, float *out_t, float *bary)
{
    v3 A = {t[0] - org.x, t[1] - org.y, t[2] - org.z};
    v3 B = {t[3] - org.x, t[4] - org.y, t[5] - org.z};
    v3 C = {t[6] - org.x, t[7] - org.y, t[8] - org.z};
    float a_kz = cmp(A, R.kz), a_kx = cmp(A, R.kx), a_ky = cmp(A, R.ky);
    float b_kz = cmp(B, R.kz), b_kx = cmp(B, R.kx), b_ky = cmp(B, R.ky);
    float c_kz = cmp(C, R.kz), c_kx = cmp(C, R.kx), c_ky = cmp(C, R.ky);
    float Ax = fmaf(-R.Sx, a_kz, a_kx), Ay = fmaf(-R.Sy, a_kz, a_ky);
    float Bx = fmaf(-R.Sx, b_kz, b_kx), By = fmaf(-R.Sy, b_kz, b_ky);
    float Cx = fmaf(-R.Sx, c_kz, c_kx), Cy = fmaf(-R.Sy, c_kz, c_ky);
    float U = fmaf(Cx, By, -Cy * Bx);
    float V = fmaf(Ax, Cy, -Ay * Cx);
    float W = fmaf(Bx, Ay, -By * Ax);
    const float eps = 1e-7f;
    if (U > -eps && U < eps) U = 0.f;
    if (V > -eps && V < eps) V = 0.f;
    if (W > -eps && W < eps) W = 0.f;
    if (U < 0.f || V < 0.f || W < 0.f) return 0;
    if (U == 0.f || V == 0.f || W == 0.f) {
        U = (float)((double)Cx * (double)By - (double)Cy * (double)Bx);
        V = (float)((double)Ax * (double)Cy - (double)Ay * (double)Cx);
        W = (float)((double)Bx * (double)Ay - (double)By * (double)Ax);
        if (U < 0.f || V < 0.f || W < 0.f) return 0;
    }
    float det = U + V + W;
    if (det == 0.f) return 0;
    float Az = R.Sz * a_kz, Bz = R.Sz * b_kz, Cz = R.Sz * c_kz;
    float T = fmaf(U, Az, fmaf(V, Bz, W * Cz));
    if (T < 0.f || T > t_max * det) return 0;
    const float rcp = 1.0f / det;
    *out_t = T * rcp;
    bary[0] = U * rcp; bary[1] = V * rcp; bary[2] = W * rcp;
    return 1;
}

/* Optional texturing state (set once by the test): per-triangle texture coordinates, the
 * texture index of each (object, submesh) material (-1 none) and up to 3 RGBA8 images.
 * Sampling restates what `tex2D<float4>` does for the reference's texture objects
 * (M/src/render/asset_processor.cpp:312-349: wrap addressing, linear filter, normalised
 * coordinates, normalised-float reads) per the CUDA programming guide's texture-fetching
 * appendix: xB = frac(u) * W - 0.5, i = floor(xB), blend weight = frac(xB) held with 8
 * fractional bits; material lookup as bvh_raycast.cpp:783-797. */
static const float *g_tri_uv = 0;
static const int *g_sub_tex = 0;
static const uint8_t *g_tex[3] = {0, 0, 0};
static int g_tex_w[3], g_tex_h[3];

/* Optional reference tracer: fn(mesh, o, d, t_max, out {t, normal xyz, uv}, &tri) -> hit; one
 * mesh handle per object id (NULL = no mesh).  The handles hold the same triangles as `tris`. */
typedef int (*hs_trace_fn)(void *mesh, const float *o, const float *d, float t_max, float *out, int *tri);
static hs_trace_fn g_trace = 0;
static void *g_mesh[8] = {0, 0, 0, 0, 0, 0, 0, 0};
void hs_oracle_set_tracer(hs_trace_fn fn, void **meshes)
{
    g_trace = fn;
    for (int i = 0; i < 8; i++) g_mesh[i] = (fn && meshes) ? meshes[i] : 0;
}

void hs_oracle_set_uv(const float *tri_uv, const int *sub_tex) { g_tri_uv = tri_uv; g_sub_tex = sub_tex; }
void hs_oracle_set_texture(int idx, const uint8_t *rgba, int w, int h) { g_tex[idx] = rgba; g_tex_w[idx] = w; g_tex_h[idx] = h; }

static int wrapi(int i, int n) { i %= n; return i < 0 ? i + n : i; }

static void sample_tex(int idx, float u, float v, float out[3])
{
    const int W = g_tex_w[idx], H = g_tex_h[idx];
    float fu = u - floorf(u), fv = v - floorf(v);
    float xb = fu * (float)W - 0.5f, yb = fv * (float)H - 0.5f;
    float xf = floorf(xb), yf = floorf(yb);
    float a = floorf((xb - xf) * 256.f + 0.5f) / 256.f, b = floorf((yb - yf) * 256.f + 0.5f) / 256.f;
    int i0 = wrapi((int)xf, W), i1 = wrapi((int)xf + 1, W), j0 = wrapi((int)yf, H), j1 = wrapi((int)yf + 1, H);
    for (int k = 0; k < 3; k++) {
        float t00 = g_tex[idx][((size_t)j0 * W + i0) * 4 + k] / 255.f, t10 = g_tex[idx][((size_t)j0 * W + i1) * 4 + k] / 255.f;
        float t01 = g_tex[idx][((size_t)j1 * W + i0) * 4 + k] / 255.f, t11 = g_tex[idx][((size_t)j1 * W + i1) * 4 + k] / 255.f;
        out[k] = (1 - a) * (1 - b) * t00 + a * (1 - b) * t10 + (1 - a) * b * t01 + a * b * t11;
    }
}

/* Renders one view.  Triangles are given per object id: object o owns triangles
 * [obj_first_tri[o], obj_first_tri[o] + obj_num_tri[o]) of `tris` ([*][9] floats, object
 * space); tri_sub[t] = submesh (material slot) of triangle t; sub_color = [num_obj][3][3].
 * Instances: object id, position, rotation (w,x,y,z), scale.  Returns 0. */
int hs_oracle_render_view(const float *tris, const int *tri_sub, const int *obj_first_tri,
                          const int *obj_num_tri, const float *sub_color,
                          int num_inst, const int *inst_obj, const float *inst_pos,
                          const float *inst_rot, const float *inst_scale,
                          const float *cam_pos, const float *cam_rot, float fov_deg,
                          int width, int height, float *depth_out, uint8_t *rgb_out)
{
    quat cq = {cam_rot[0], cam_rot[1], cam_rot[2], cam_rot[3]};
    v3 start = {cam_pos[0], cam_pos[1], cam_pos[2] + 0.5f};
    const float to_rad = 3.14159265358979323846f / 180.f;
    float fov_scale = 1.0f / tanf(to_rad * (fov_deg * 0.5f));
    float y_scale = -fov_scale;
    v3 fy = {0, 1, 0}, fx = {1, 0, 0};
    v3 look_at = rot(cq, fy);
    float h = 1.0f / (-y_scale);
    float vp = 2 * h;
    v3 forward = nrm(look_at);
    v3 u = rot(cq, fx);
    v3 v = nrm(crs(forward, u));
    v3 horizontal = scl(u, vp), vertical = scl(v, vp);
    v3 ll = add(sub(sub(start, scl(horizontal, 1.f / 2.f)), scl(vertical, 1.f / 2.f)), forward);

    for (int py = 0; py < height; py++) {
        for (int px = 0; px < width; px++) {
            float pu = ((float)px + 0.5f) / (float)width;
            float pv = ((float)py + 0.5f) / (float)height;
            v3 d = nrm(sub(add(add(ll, scl(horizontal, pu)), scl(vertical, pv)), start));
            float t_max = 10000.f;
            int hit = 0, hit_obj = 0, hit_sub = 0, hit_tri = 0;
            float hit_uv[2] = {0, 0};
            for (int i = 0; i < num_inst; i++) {
                int o = inst_obj[i];
                const float *s = inst_scale + 3 * i;
                if (s[0] == 0.f && s[1] == 0.f && s[2] == 0.f) continue;
                if (obj_num_tri[o] == 0) continue;
                quat qi = qconj((quat){inst_rot[4 * i], inst_rot[4 * i + 1], inst_rot[4 * i + 2], inst_rot[4 * i + 3]});
                v3 p = {inst_pos[3 * i], inst_pos[3 * i + 1], inst_pos[3 * i + 2]};
                v3 ro = rot(qi, sub(start, p));
                ro.x *= 1.f / s[0]; ro.y *= 1.f / s[1]; ro.z *= 1.f / s[2];
                v3 rd = rot(qi, d);
                rd.x *= 1.f / s[0]; rd.y *= 1.f / s[1]; rd.z *= 1.f / s[2];
                float t_scale = len(rd);
                float tm = t_max * t_scale;
                rd = scl(rd, 1.f / t_scale);
                int any = 0;
                if (g_trace && g_mesh[o]) {
                    /* the reference's MeshBVH::traceRay on this instance's mesh */
                    float out[6];
                    int tri = 0;
                    float rof[3] = {ro.x, ro.y, ro.z}, rdf[3] = {rd.x, rd.y, rd.z};
                    if (g_trace(g_mesh[o], rof, rdf, tm, out, &tri)) {
                        tm = out[0]; any = 1; hit_tri = obj_first_tri[o] + tri; hit_sub = tri_sub[hit_tri];
                        hit_uv[0] = out[4]; hit_uv[1] = out[5];
                    }
                } else {
                    isect R = isect_info(rd);
                    for (int t = 0; t < obj_num_tri[o]; t++) {
                        int ti = obj_first_tri[o] + t;
                        float th, bc[3];
                        if (tri_hit(tris + 9 * ti, R, ro, tm, &th, bc)) {
                            tm = th; any = 1; hit_sub = tri_sub[ti]; hit_tri = ti;
                            if (g_tri_uv) {
                                const float *uv = g_tri_uv + 6 * ti;
                                hit_uv[0] = uv[0] * bc[0] + uv[2] * bc[1] + uv[4] * bc[2];
                                hit_uv[1] = uv[1] * bc[0] + uv[3] * bc[1] + uv[5] * bc[2];
                            }
                        }
                    }
                }
                if (any) { t_max = tm / t_scale; hit = 1; hit_obj = o; }
            }
            size_t pix = (size_t)py * width + px;
            uint8_t *rgb = rgb_out + 4 * pix;
            if (hit) {
                const float *mc = sub_color + (hit_obj * 3 + hit_sub) * 3;
                float c[3] = {mc[0], mc[1], mc[2]};
                int tex = g_sub_tex ? g_sub_tex[hit_obj * 3 + hit_sub] : -1;
                if (tex >= 0 && g_tex[tex] && g_tri_uv) {
                    float tu = hit_uv[0], tv = hit_uv[1];
                    float tc[3];
                    sample_tex(tex, tu, 1.f - tv, tc);
                    for (int k = 0; k < 3; k++) c[k] *= tc[k];
                }
                depth_out[pix] = t_max;
                for (int k = 0; k < 3; k++) rgb[k] = (uint8_t)(fminf(1.f, fmaxf(0.2f, 0.f) * c[k]) * 255);
            } else {
                depth_out[pix] = 0.f;
                rgb[0] = rgb[1] = rgb[2] = 0;
            }
            rgb[3] = 255;
        }
    }
    return 0;
}

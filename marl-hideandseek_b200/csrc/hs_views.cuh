// Agent camera views: per-agent depth + RGB images ray-cast over the world's render instances.
//
// Replaces the reference's batch renderer output (Manager::depthTensor / rgbTensor,
// src/mgr.cpp:1241-1263) with a CUDA ray caster that follows the arithmetic of Madrona's
// software ray caster (M/src/mw/device/bvh_raycast.cpp):
//   * camera: entity position + 0.5*up, entity rotation, vertical fov 100 deg
//     (src/sim.cpp:1399-1403; viewTransformUpdate M/src/render/ecs_system.cpp:274-313);
//   * primary ray per pixel: calculateOutRay (bvh_raycast.cpp:58-88);
//   * per instance the ray is moved into object space, re-normalised, and tested against the
//     render mesh with the watertight test of Woop et al. with back-face culling
//     (bvh_raycast.cpp:228-270, 317-447, 614-633; MADRONA_MESHBVH_BACKFACE_CULLING is defined
//     in M/include/madrona/mesh_bvh.inl:3);
//   * depth = distance along the unit ray to the closest hit, 0 on a miss; this environment
//     creates no light entities, so computeFragment's colour is max(0.2, 0) * material colour
//     (bvh_raycast.cpp:848-935); rgb bytes = colour * 255 truncated, alpha 255 (:822-839).
// What differs by design: no TLAS/BLAS walk with a per-thread stack.  Persistent CTAs keep the
// whole render-mesh table in shared memory; per view, the world's <= 52 instances are staged in
// shared memory front to back with their object-space maps; which triangles (ground / boxes /
// wedge) or triangle clusters (agents) face the camera is decided once per view, not per pixel;
// a warp culls the instance list against its 8x4-pixel tile's pyramid cooperatively, and each
// ray then walks the survivors (slab test against the closest hit so far, then triangles).
// This file is its own translation unit (hs_views.cu) compiled with FMA contraction and fast
// division: nothing here has to match a CPU bit for bit.
// The three textures (ground grid, hider / seeker faces) are sampled through CUDA texture
// objects set up like the reference's when a data directory is given; the built-in meshes have
// no texture coordinates and use the material's base colour.  Inactive agent slots produce an
// all-zero view.
#pragma once

#include "hs_state.cuh"

namespace hs {

constexpr int kViewMaxTris = 1552;
constexpr int kViewMaxClusters = 200;
constexpr int kViewMaxSubmeshes = 16;
constexpr int kViewMaxSubPerObj = 3;
constexpr int kViewClusterTris = 8;       // <= 64 clusters per submesh (one 64-bit mask)
#ifndef HS_VIEW_THREADS
#define HS_VIEW_THREADS 256
#endif
#ifndef HS_VIEW_CTAS
#define HS_VIEW_CTAS 3
#endif
#ifndef HS_VIEW_TILE_W
#define HS_VIEW_TILE_W 8
#endif
constexpr int kViewThreads = HS_VIEW_THREADS;
constexpr int kViewTileW = HS_VIEW_TILE_W, kViewTileH = 32 / HS_VIEW_TILE_W;   // a warp's pixel tile
constexpr int kViewSmallTris = 16;     // a mesh of at most this many triangles is one flat list with a 16-bit facing mask
constexpr int kViewCtasPerSm = HS_VIEW_CTAS;

struct ViewBox { float lo[3]; int first; float hi[3]; int count; };     // 32 B
struct ViewMeshTable {
    int objFirstSub[OBJ_COUNT], objNumSub[OBJ_COUNT];
    float objAabb[OBJ_COUNT][6];
    float subColor[OBJ_COUNT][kViewMaxSubPerObj][3];   // material base colour of (object, submesh)
    ViewBox sub[kViewMaxSubmeshes];                    // first/count index clusters
    ViewBox cluster[kViewMaxClusters];                 // first/count index triangles
    int numSub, numClusters, numTris;
    int numSmallTris;                                  // tri[0, numSmallTris): the single-cluster meshes, stored first
    float fovScale;                                    // 1 / tanf(toRadians(100 / 2))
    float tri[kViewMaxTris][9];                        // a, b, c in object space
};

// What only the final shading of a pixel needs (stays in global memory): texture coordinates of
// every triangle, and which texture a (object, submesh) material samples (-1: base colour only).
constexpr int kViewTextures = 3;
struct ViewTexTable {
    float triUV[kViewMaxTris][6];
    int subTex[OBJ_COUNT][kViewMaxSubPerObj];
    unsigned long long tex[kViewTextures];             // cudaTextureObject_t: RGBA8, wrap, linear, normalised
};

struct ViewParams {
    const ViewMeshTable *mesh;
    const ViewTexTable *texTable;
    float *depth;            // [N*A, H, W, 1]
    uint8_t *rgb;            // [N*A, H, W, 4]
    int width, height;
};

#if defined(__CUDACC__) && !defined(HS_VIEWS_HOST_ONLY)

struct ViewInst {
    Vec3 lo; int obj;        // world AABB of the instance's render mesh (padded); obj -1: nothing to
    Vec3 hi; uint32_t mask;  //   hit (no mesh, zero scale, or no camera-facing triangle)
    Vec3 ro;                 // camera origin in object space
    // world -> object direction map scale^-1 * R^T, by columns (images of the world axes)
    Vec3 cx, cy, cz;
    // single-cluster meshes (ground, boxes, wedge): triangle range in the shared-memory copy;
    // `mask` = the subset that faces the camera.  triCount == 0 marks a clustered mesh.
    int triFirst, triCount;
    // clustered meshes (agents): per submesh, the clusters holding a triangle that faces the camera
    unsigned long long cmask[kViewMaxSubPerObj];
};

struct ViewSmem {
    ViewInst inst[kMaxBodies];           // front to back by distance from the camera to the box
    float key[kMaxBodies];
    int active;
    int nextTile;                        // tiles are handed to the warps dynamically (their cost varies a lot)
    Vec3 origin, lowerLeft, horizontal, vertical;
};

// Woop et al. ray constants: permutation (kz = dominant axis), shear, and the ray origin permuted
struct WoopRay { int kx, ky, kz; float Sx, Sy, Sz; float ox, oy, oz; };

// computeRayIsectInfo (bvh_raycast.cpp:228-270)
__device__ __forceinline__ WoopRay woopSetup(Vec3 d, Vec3 org)
{
    float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    int kz = (ax > ay && ax > az) ? 0 : (ay > az ? 1 : 2);
    int kx = kz + 1; if (kx == 3) kx = 0;
    int ky = kx + 1; if (ky == 3) ky = 0;
    float dz = comp(d, kz);
    if (dz < 0.f) { int t = kx; kx = ky; ky = t; }
    float inv = 1.f / dz;
    return {kx, ky, kz, comp(d, kx) * inv, comp(d, ky) * inv, inv, comp(org, kx), comp(org, ky), comp(org, kz)};
}

// rayTriangleIntersection with back-face culling (bvh_raycast.cpp:317-447); returns the hit
// distance in object-space units of the normalised ray, or a negative value on a miss.  The
// axis permutation is applied by indexing the triangle's nine floats, not by selecting
// between loaded components.
__device__ __forceinline__ float woopHit(const float *t9, const WoopRay &R, float t_max)
{
    const float a_kx = t9[R.kx] - R.ox, a_ky = t9[R.ky] - R.oy, a_kz = t9[R.kz] - R.oz;
    const float b_kx = t9[3 + R.kx] - R.ox, b_ky = t9[3 + R.ky] - R.oy, b_kz = t9[3 + R.kz] - R.oz;
    const float c_kx = t9[6 + R.kx] - R.ox, c_ky = t9[6 + R.ky] - R.oy, c_kz = t9[6 + R.kz] - R.oz;
    float Ax = fmaf(-R.Sx, a_kz, a_kx), Ay = fmaf(-R.Sy, a_kz, a_ky);
    float Bx = fmaf(-R.Sx, b_kz, b_kx), By = fmaf(-R.Sy, b_kz, b_ky);
    float Cx = fmaf(-R.Sx, c_kz, c_kx), Cy = fmaf(-R.Sy, c_kz, c_ky);
    float U = fmaf(Cx, By, -Cy * Bx);
    float V = fmaf(Ax, Cy, -Ay * Cx);
    float W = fmaf(Bx, Ay, -By * Ax);
    const float eps = 1e-7f;
    if (fabsf(U) < eps) U = 0.f;
    if (fabsf(V) < eps) V = 0.f;
    if (fabsf(W) < eps) W = 0.f;
    if (fminf(U, fminf(V, W)) < 0.f) return -1.f;
    if (U == 0.f || V == 0.f || W == 0.f) {
        U = (float)((double)Cx * (double)By - (double)Cy * (double)Bx);
        V = (float)((double)Ax * (double)Cy - (double)Ay * (double)Cx);
        W = (float)((double)Bx * (double)Ay - (double)By * (double)Ax);
        if (U < 0.f || V < 0.f || W < 0.f) return -1.f;
    }
    float det = U + V + W;
    if (det == 0.f) return -1.f;
    float Az = R.Sz * a_kz, Bz = R.Sz * b_kz, Cz = R.Sz * c_kz;
    float T = fmaf(U, Az, fmaf(V, Bz, W * Cz));
    if (T < 0.f || T > t_max * det) return -1.f;
    return T * (1.f / det);
}

// Normalised barycentrics (U, V, W) / det of a triangle the ray is known to hit — the same
// arithmetic as woopHit up to the determinant (bvh_raycast.cpp:331-392, 436-439); used once per
// pixel to interpolate the texture coordinates of the final hit (triangleIntersect :474-476).
__device__ __forceinline__ Vec3 woopBary(const float *t9, const WoopRay &R)
{
    const float a_kx = t9[R.kx] - R.ox, a_ky = t9[R.ky] - R.oy, a_kz = t9[R.kz] - R.oz;
    const float b_kx = t9[3 + R.kx] - R.ox, b_ky = t9[3 + R.ky] - R.oy, b_kz = t9[3 + R.kz] - R.oz;
    const float c_kx = t9[6 + R.kx] - R.ox, c_ky = t9[6 + R.ky] - R.oy, c_kz = t9[6 + R.kz] - R.oz;
    float Ax = fmaf(-R.Sx, a_kz, a_kx), Ay = fmaf(-R.Sy, a_kz, a_ky);
    float Bx = fmaf(-R.Sx, b_kz, b_kx), By = fmaf(-R.Sy, b_kz, b_ky);
    float Cx = fmaf(-R.Sx, c_kz, c_kx), Cy = fmaf(-R.Sy, c_kz, c_ky);
    float U = fmaf(Cx, By, -Cy * Bx);
    float V = fmaf(Ax, Cy, -Ay * Cx);
    float W = fmaf(Bx, Ay, -By * Ax);
    const float eps = 1e-7f;
    if (fabsf(U) < eps) U = 0.f;
    if (fabsf(V) < eps) V = 0.f;
    if (fabsf(W) < eps) W = 0.f;
    if (U == 0.f || V == 0.f || W == 0.f) {
        U = (float)((double)Cx * (double)By - (double)Cy * (double)Bx);
        V = (float)((double)Ax * (double)Cy - (double)Ay * (double)Cx);
        W = (float)((double)Bx * (double)Ay - (double)By * (double)Ax);
    }
    const float rcp = 1.f / (U + V + W);
    return {U * rcp, V * rcp, W * rcp};
}

__device__ __forceinline__ bool viewSlab(const float *lo, const float *hi, Vec3 o, Vec3 inv_d, float t_max)
{
    return aabbRayIntersects({lo[0], lo[1], lo[2]}, {hi[0], hi[1], hi[2]}, o, inv_d, 0.f, t_max);
}

// A triangle can only be hit from its front side (back-face culling), and which side the camera
// is on does not depend on the pixel: decide it once per (view, instance).  The test keeps a
// triangle whenever the camera is not clearly behind it, so it never drops one the per-ray
// test would have accepted.
__device__ __forceinline__ bool viewFacesCamera(const float *t9, Vec3 ro)
{
    Vec3 a = {t9[0], t9[1], t9[2]};
    Vec3 n = cross(v3(t9[3], t9[4], t9[5]) - a, v3(t9[6], t9[7], t9[8]) - a);
    Vec3 r = ro - a;
    return dot(n, r) >= -1e-4f * length(n) * fmaxf(1.f, length(r));
}

// Is the box [lo,hi] (relative to the apex) entirely outside the half space n.x >= 0 ?
__device__ __forceinline__ bool viewBoxBehind(Vec3 n, Vec3 lo, Vec3 hi)
{
    float m = fmaxf(n.x * lo.x, n.x * hi.x) + fmaxf(n.y * lo.y, n.y * hi.y) + fmaxf(n.z * lo.z, n.z * hi.z);
    return m < 0.f;
}

// Persistent CTAs (kViewCtasPerSm per SM): the whole render-mesh table (64 KB) is copied into
// shared memory once per CTA, then the CTA takes views (world, agent slot) round-robin; a warp
// shades 8x4 pixel tiles.
__global__ void __launch_bounds__(kViewThreads, kViewCtasPerSm) k_agent_views(DeviceState S, ViewParams P, int numViews)
{
    __shared__ ViewSmem sm;
    extern __shared__ uint4 viewMeshSmem[];
    const ViewMeshTable &M = *reinterpret_cast<const ViewMeshTable *>(viewMeshSmem);
    const int tid = threadIdx.x;
    {
        const uint4 *src = reinterpret_cast<const uint4 *>(P.mesh);
        for (int i = tid; i < int(sizeof(ViewMeshTable) / sizeof(uint4)); i += kViewThreads) viewMeshSmem[i] = src[i];
    }
    const int A = S.p.agentsPerWorld;
    const int W = P.width, H = P.height;
  for (int view = blockIdx.x; view < numViews; view += gridDim.x) {
    const int w = view / A, k = view - w * A;
    __syncthreads();         // the table copy (first round) / the previous view's readers of sm

    if (tid == 0) {
        int slot = WB(S, WB_IFACE_SLOT + k, w);
        sm.active = slot != 0xFF;
        sm.nextTile = 0;
        if (sm.active) {
            // viewTransformUpdate + calculateOutRay
            Pose ps = loadPose(S, REC_POSE, slot, w);
            Vec3 start = ps.x + v3(0.f, 0.f, 0.5f);
            Vec3 forward = normalize(rotate(ps.q, v3(0.f, 1.f, 0.f)));
            Vec3 u = rotate(ps.q, v3(1.f, 0.f, 0.f));
            Vec3 v = normalize(cross(forward, u));
            float h = 1.f / M.fovScale;
            float vp = 2.f * h;
            sm.horizontal = u * vp;
            sm.vertical = v * vp;
            sm.origin = start;
            sm.lowerLeft = start - sm.horizontal / 2.f - sm.vertical / 2.f + forward;
        }
    }
    __syncthreads();
    const size_t pixBase = size_t(view) * W * H;
    if (!sm.active) {
        for (int i = tid; i < W * H; i += kViewThreads) {
            P.depth[pixBase + i] = 0.f;
            reinterpret_cast<uint32_t *>(P.rgb)[pixBase + i] = 0u;
        }
        continue;
    }
    const int nBodies = WI(S, WF_NUM_BODIES, w);
    ViewInst I;
    float key = 0.f;
    bool skip = true;
    if (tid < nBodies) {
        uint32_t m = META(S, tid, w);
        I.obj = metaObj(m);
        Pose ps = loadPose(S, REC_POSE, tid, w);
        Vec3 sc = loadScale(S, tid, w);
        // instanceHasVolume (bvh_raycast.cpp:272-277) and meshes this table does not carry
        skip = (sc.x == 0.f && sc.y == 0.f && sc.z == 0.f) || M.objNumSub[I.obj] == 0;
        const int obj = skip ? 0 : I.obj;
        const float *bb = M.objAabb[obj];
        AABB wb = aabbApplyTRS({{bb[0], bb[1], bb[2]}, {bb[3], bb[4], bb[5]}}, ps.x, ps.q, sc);
        // the cull must never reject a ray the triangle test would accept: pad by rounding slack
        Vec3 ext = (wb.hi - wb.lo) * 1e-5f + v3(1e-5f, 1e-5f, 1e-5f);
        I.lo = wb.lo - ext; I.hi = wb.hi + ext;
        // ray into object space (bvh_raycast.cpp:625-633): scale^-1 * rotateVec(rot^-1, .)
        const Quat qi = qinv(ps.q);
        const Vec3 invS = {1.f / sc.x, 1.f / sc.y, 1.f / sc.z};
        I.ro = mulDiag(invS, rotate(qi, sm.origin - ps.x));
        I.cx = mulDiag(invS, rotate(qi, v3(1.f, 0.f, 0.f)));
        I.cy = mulDiag(invS, rotate(qi, v3(0.f, 1.f, 0.f)));
        I.cz = mulDiag(invS, rotate(qi, v3(0.f, 0.f, 1.f)));
        I.triFirst = 0; I.triCount = 0; I.mask = 0;
        for (int s = 0; s < kViewMaxSubPerObj; s++) I.cmask[s] = 0;
        if (!skip && M.objNumSub[obj] == 1) {
            const ViewBox &SB = M.sub[M.objFirstSub[obj]];
            const ViewBox &CB = M.cluster[SB.first];
            if (SB.count == 1 && CB.count <= kViewSmallTris) {
                I.triFirst = CB.first; I.triCount = CB.count;
                for (int t = 0; t < CB.count; t++)
                    if (viewFacesCamera(M.tri[CB.first + t], I.ro)) I.mask |= 1u << t;
                if (I.mask == 0) skip = true;
            }
        }
        if (skip) I.obj = -1;
        // distance from the camera to the box (0 inside): the visiting order
        Vec3 o = sm.origin;
        Vec3 dl = vmax(vmax(I.lo - o, o - I.hi), v3(0.f, 0.f, 0.f));
        key = skip ? 3.0e38f : length2(dl);
        sm.key[tid] = key;
    }
    const int live = __syncthreads_count(!skip);
    if (tid < nBodies) {
        int rank = 0;
        for (int j = 0; j < nBodies; j++) {
            float kj = sm.key[j];
            rank += (kj < key || (kj == key && j < tid)) ? 1 : 0;
        }
        sm.inst[rank] = I;
    }
    __syncthreads();
    const int nInst = live;
    // clustered instances: thread per cluster marks the clusters with a camera-facing triangle
    // (a camera inside a closed mesh — every agent sits inside its own body — faces none of it)
    for (int i = 0; i < nInst; i++) {
        if (sm.inst[i].triCount > 0) continue;
        const int obj = sm.inst[i].obj;
        const int s0 = M.objFirstSub[obj], ns = M.objNumSub[obj];
        const Vec3 ro = sm.inst[i].ro;
        for (int s = 0; s < ns; s++) {
            const ViewBox &SB = M.sub[s0 + s];
            for (int c = tid; c < SB.count; c += kViewThreads) {
                const ViewBox &CB = M.cluster[SB.first + c];
                bool any = false;
                for (int t = 0; t < CB.count; t++) any |= viewFacesCamera(M.tri[CB.first + t], ro);
                if (any) atomicOr(&sm.inst[i].cmask[s], 1ull << c);
            }
        }
    }
    __syncthreads();

    const int lane = tid & 31;
    const int tilesX = (W + kViewTileW - 1) / kViewTileW, tilesY = (H + kViewTileH - 1) / kViewTileH;
    const Vec3 o = sm.origin;
    const Vec3 LL = sm.lowerLeft - o, HZ = sm.horizontal, VT = sm.vertical;
    const float sU = dot(cross(LL, VT), HZ) < 0.f ? -1.f : 1.f, sV = dot(cross(LL, HZ), VT) < 0.f ? -1.f : 1.f;
    for (;;) {
        int tile = 0;
        if (lane == 0) tile = atomicAdd(&sm.nextTile, 1);
        tile = __shfl_sync(0xffffffffu, tile, 0);
        if (tile >= tilesX * tilesY) break;
        const int px0 = (tile % tilesX) * kViewTileW, py0 = (tile / tilesX) * kViewTileH;
        // Warp-cooperative cull: lane j tests instances j and j+32 against the pyramid of this
        // 8x4-pixel tile (apex = camera, four side planes through the tile's corner rays); the
        // rays then only visit the instances some ray of the tile can reach, still front to back.
        uint64_t todo;
        {
            const float u0 = float(px0) / float(W), u1 = float(min(px0 + kViewTileW, W)) / float(W);
            const float v0 = float(py0) / float(H), v1 = float(min(py0 + kViewTileH, H)) / float(H);
            // side planes of the pyramid, normals pointing inward.  All directions of one pixel
            // column LL + u*HZ + v*VT (u fixed) lie in the plane with normal (LL + u*HZ) x VT, so
            // a plane needs one cross product; which way is "inward" is the same for every tile.
            const Vec3 nl = sU * cross(LL + u0 * HZ, VT), nr = -sU * cross(LL + u1 * HZ, VT);
            const Vec3 nt = sV * cross(LL + v0 * VT, HZ), nb = -sV * cross(LL + v1 * VT, HZ);
            bool keep[2];
#pragma unroll
            for (int r = 0; r < 2; r++) {
                const int j = lane + 32 * r;
                keep[r] = false;
                if (j < nInst) {
                    // widen the box a little: the planes come out of rounded corner directions
                    const Vec3 lo = sm.inst[j].lo - o, hi = sm.inst[j].hi - o;
                    const Vec3 e = (hi - lo) * 1e-4f + v3(1e-4f, 1e-4f, 1e-4f);
                    const Vec3 l2 = lo - e, h2 = hi + e;
                    keep[r] = !(viewBoxBehind(nl, l2, h2) || viewBoxBehind(nr, l2, h2) ||
                                viewBoxBehind(nt, l2, h2) || viewBoxBehind(nb, l2, h2));
                }
            }
            todo = uint64_t(__ballot_sync(0xffffffffu, keep[0])) | (uint64_t(__ballot_sync(0xffffffffu, keep[1])) << 32);
        }
        const int px = px0 + (lane % kViewTileW), py = py0 + (lane / kViewTileW);
        // lanes of a partial edge tile that fall outside the image keep walking with the warp (every
        // lane meets the warp-wide shuffle / ballots at the loop top together); they trace nothing
        const bool inImage = px < W && py < H;
        if (!inImage) todo = 0;
        float pu = (float(px) + 0.5f) / float(W);
        float pv = (float(py) + 0.5f) / float(H);
        Vec3 d = normalize(sm.lowerLeft + pu * sm.horizontal + pv * sm.vertical - o);
        Vec3 invd = {1.f / d.x, 1.f / d.y, 1.f / d.z};
        float best = 10000.f;
        int hitObj = -1, hitSub = 0, hitInst = 0, hitTri = 0;
        for (; todo; todo &= todo - 1) {
            const int i = __ffsll((long long)todo) - 1;
            if (!aabbRayIntersects(sm.inst[i].lo, sm.inst[i].hi, o, invd, 0.f, best)) continue;
            Vec3 rd = d.x * sm.inst[i].cx + d.y * sm.inst[i].cy + d.z * sm.inst[i].cz;
            float ts = length(rd);
            rd = rd / ts;
            const Vec3 ro = sm.inst[i].ro;
            float tmax = best * ts;
            const WoopRay R = woopSetup(rd, ro);
            bool hit = false;
            int tri = 0;
            const int triCount = sm.inst[i].triCount;
            if (triCount > 0) {
                const int first = sm.inst[i].triFirst;
                for (uint32_t m = sm.inst[i].mask; m; m &= m - 1) {
                    const int t = first + __ffs(m) - 1;
                    float th = woopHit(M.tri[t], R, tmax);
                    if (th >= 0.f) { tmax = th; hit = true; hitSub = 0; tri = t; }
                }
            } else {
                Vec3 invrd = {1.f / rd.x, 1.f / rd.y, 1.f / rd.z};
                const int obj = sm.inst[i].obj;
                const int s0 = M.objFirstSub[obj], ns = M.objNumSub[obj];
                for (int s = 0; s < ns; s++) {
                    unsigned long long cm = sm.inst[i].cmask[s];
                    if (cm == 0) continue;
                    const ViewBox &SB = M.sub[s0 + s];
                    if (ns > 1 && !viewSlab(SB.lo, SB.hi, ro, invrd, tmax)) continue;
                    for (; cm; cm &= cm - 1) {
                        const ViewBox &CB = M.cluster[SB.first + __ffsll((long long)cm) - 1];
                        if (!viewSlab(CB.lo, CB.hi, ro, invrd, tmax)) continue;
                        for (int t = 0; t < CB.count; t++) {
                            float th = woopHit(M.tri[CB.first + t], R, tmax);
                            if (th >= 0.f) { tmax = th; hit = true; hitSub = s; tri = CB.first + t; }
                        }
                    }
                }
            }
            if (hit) { best = tmax / ts; hitObj = sm.inst[i].obj; hitInst = i; hitTri = tri; }
        }
        const size_t pix = pixBase + size_t(py) * W + px;
        uint32_t rgba = 0xFF000000u;
        float depth = 0.f;
        if (hitObj >= 0) {
            depth = best;
            const float *col = M.subColor[hitObj][hitSub];
            Vec3 c = {col[0], col[1], col[2]};
            const int texId = P.texTable->subTex[hitObj][hitSub];
            if (texId >= 0) {
                // traceRay's material lookup (bvh_raycast.cpp:783-797): interpolate the hit
                // triangle's texture coordinates, sample at (u, 1 - v), modulate the base colour
                Vec3 rd = d.x * sm.inst[hitInst].cx + d.y * sm.inst[hitInst].cy + d.z * sm.inst[hitInst].cz;
                rd = rd / length(rd);
                const Vec3 bary = woopBary(M.tri[hitTri], woopSetup(rd, sm.inst[hitInst].ro));
                const float *uv = P.texTable->triUV[hitTri];
                const float tu = uv[0] * bary.x + uv[2] * bary.y + uv[4] * bary.z;
                const float tv = uv[1] * bary.x + uv[3] * bary.y + uv[5] * bary.z;
                const float4 tc = tex2D<float4>((cudaTextureObject_t)P.texTable->tex[texId], tu, 1.f - tv);
                c = {tc.x * c.x, tc.y * c.y, tc.z * c.z};
            }
            // acc_color = fmaxf(0.2, light_contrib = 0) * color, clamped to 1; bytes truncate
            uint32_t r = (uint32_t)(fminf(1.f, 0.2f * c.x) * 255.f);
            uint32_t g = (uint32_t)(fminf(1.f, 0.2f * c.y) * 255.f);
            uint32_t b = (uint32_t)(fminf(1.f, 0.2f * c.z) * 255.f);
            rgba |= r | (g << 8) | (b << 16);
        }
        if (inImage) {
            P.depth[pix] = depth;
            reinterpret_cast<uint32_t *>(P.rgb)[pix] = rgba;
        }
    }
  }
}

#endif  // __CUDACC__

// host side (hs_view_assets.cpp): fills the table from the reference's data directory
// (plane.obj, cube_render.obj, wall_render.obj, agent_render.obj, ramp_render.obj,
// elongated_render.obj — src/mgr.cpp:590-609) or, with dataDir == nullptr, from built-in
// descriptions of the same shapes.  Returns an empty string on success, else the error.
const char *buildViewMeshTable(ViewMeshTable &T, ViewTexTable &X, const char *dataDir);

}  // namespace hs

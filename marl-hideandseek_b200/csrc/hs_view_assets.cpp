// Init-time render-mesh table for the agent camera views (hs_views.cuh).
//
// Replaces, for this environment's fixed asset set, loadRenderObjects (src/mgr.cpp:590-672):
// the six render meshes, their per-submesh material colours, and — instead of Madrona's
// per-mesh QBVH (M/include/madrona/mesh_bvh.hpp) — a two-level box hierarchy
// (submesh -> clusters of <= 8 triangles; meshes of <= 16 triangles stay one flat list) sized for meshes of 2..1440 triangles.
//
// Geometry comes from the reference's data directory when one is given (the six OBJ files
// the reference loads); otherwise from this repo's own description of the same shapes: the
// ground quad, the four boxes and the wedge are exact, the agent's three 480-triangle round
// meshes (body + two eyes) are stood in for by frequency-5 geodesic ellipsoids with the same
// centres and radii (500 triangles each) — tests/test_views_cpu.py measures the difference
// against the reference's OBJ where that is present.
#define HS_VIEWS_HOST_ONLY
#include "hs_views.cuh"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include <zlib.h>

namespace hs {

namespace {

struct Tri { float v[9]; float uv[6]; };
struct SubSrc { std::vector<Tri> tris; };
struct MeshSrc { std::vector<SubSrc> subs; };

Tri mkTri(const float *a, const float *b, const float *c)
{
    Tri t;
    for (int k = 0; k < 3; k++) { t.v[k] = a[k]; t.v[3 + k] = b[k]; t.v[6 + k] = c[k]; }
    for (int k = 0; k < 6; k++) t.uv[k] = 0.f;
    return t;
}

// axis-aligned block, outward CCW triangles
MeshSrc block(float hx, float hy, float zb, float zt)
{
    const float P[8][3] = {{-hx, -hy, zb}, {hx, -hy, zb}, {hx, hy, zb}, {-hx, hy, zb},
                           {-hx, -hy, zt}, {hx, -hy, zt}, {hx, hy, zt}, {-hx, hy, zt}};
    const int Q[6][4] = {{0, 3, 2, 1}, {4, 5, 6, 7}, {0, 1, 5, 4}, {2, 3, 7, 6}, {1, 2, 6, 5}, {3, 0, 4, 7}};
    MeshSrc m; m.subs.resize(1);
    for (auto &q : Q) {
        m.subs[0].tris.push_back(mkTri(P[q[0]], P[q[1]], P[q[2]]));
        m.subs[0].tris.push_back(mkTri(P[q[0]], P[q[2]], P[q[3]]));
    }
    return m;
}

// the ramp: a prism over the (y,z) triangle (1,-1) (-2,-1) (1,1), x in [-1,1]
MeshSrc wedge()
{
    const float P[6][3] = {{-1, 1, -1}, {1, 1, -1}, {1, -2, -1}, {-1, -2, -1}, {-1, 1, 1}, {1, 1, 1}};
    MeshSrc m; m.subs.resize(1);
    auto quad = [&](int a, int b, int c, int d) {
        m.subs[0].tris.push_back(mkTri(P[a], P[b], P[c]));
        m.subs[0].tris.push_back(mkTri(P[a], P[c], P[d]));
    };
    quad(0, 1, 2, 3);        // bottom, faces -z
    quad(0, 4, 5, 1);        // back, faces +y
    quad(2, 5, 4, 3);        // slope
    m.subs[0].tris.push_back(mkTri(P[1], P[5], P[2]));   // +x side
    m.subs[0].tris.push_back(mkTri(P[3], P[4], P[0]));   // -x side
    return m;
}

MeshSrc groundQuad()
{
    const float P[4][3] = {{-1000, -1000, 0}, {1000, -1000, 0}, {-1000, 1000, 0}, {1000, 1000, 0}};
    MeshSrc m; m.subs.resize(1);
    m.subs[0].tris.push_back(mkTri(P[1], P[2], P[0]));
    m.subs[0].tris.push_back(mkTri(P[1], P[3], P[2]));
    return m;
}

// frequency-f geodesic ellipsoid: icosahedron with poles on z, every face cut into f*f triangles
SubSrc geodesic(int f, const float c[3], const float r[3])
{
    const double pi = 3.14159265358979323846;
    double V[12][3];
    V[0][0] = 0; V[0][1] = 0; V[0][2] = 1;
    V[11][0] = 0; V[11][1] = 0; V[11][2] = -1;
    const double zr = 1.0 / std::sqrt(5.0), rr = 2.0 / std::sqrt(5.0);
    for (int k = 0; k < 5; k++) {
        double a = 2 * pi * k / 5, b = a + pi / 5;
        V[1 + k][0] = rr * std::cos(a); V[1 + k][1] = rr * std::sin(a); V[1 + k][2] = zr;
        V[6 + k][0] = rr * std::cos(b); V[6 + k][1] = rr * std::sin(b); V[6 + k][2] = -zr;
    }
    std::vector<std::vector<int>> F;
    for (int k = 0; k < 5; k++) {
        int u0 = 1 + k, u1 = 1 + (k + 1) % 5, l0 = 6 + k, l1 = 6 + (k + 1) % 5;
        F.push_back({0, u0, u1});
        F.push_back({u0, l0, u1});
        F.push_back({u1, l0, l1});
        F.push_back({11, l1, l0});
    }
    SubSrc out;
    auto point = [&](const std::vector<int> &fc, int i, int j, float p[3]) {
        double q[3], n = 0;
        for (int k = 0; k < 3; k++) {
            q[k] = V[fc[0]][k] + (V[fc[1]][k] - V[fc[0]][k]) * i / f + (V[fc[2]][k] - V[fc[0]][k]) * j / f;
            n += q[k] * q[k];
        }
        n = std::sqrt(n);
        for (int k = 0; k < 3; k++) p[k] = float(c[k] + r[k] * q[k] / n);
    };
    auto emit = [&](const float *a, const float *b, const float *cc) {
        // orient outward
        float e1[3], e2[3], mid[3];
        for (int k = 0; k < 3; k++) { e1[k] = b[k] - a[k]; e2[k] = cc[k] - a[k]; mid[k] = (a[k] + b[k] + cc[k]) / 3.f - c[k]; }
        float n[3] = {e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0]};
        if (n[0] * mid[0] + n[1] * mid[1] + n[2] * mid[2] >= 0) out.tris.push_back(mkTri(a, b, cc));
        else out.tris.push_back(mkTri(a, cc, b));
    };
    for (auto &fc : F)
        for (int i = 0; i < f; i++)
            for (int j = 0; i + j < f; j++) {
                float a[3], b[3], cc[3], d[3];
                point(fc, i, j, a); point(fc, i + 1, j, b); point(fc, i, j + 1, cc);
                emit(a, b, cc);
                if (i + j < f - 1) { point(fc, i + 1, j + 1, d); emit(b, d, cc); }
            }
    return out;
}

MeshSrc agentStandIn()
{
    // centres / radii of the three round meshes of the reference's agent render asset
    const float bc[3] = {0.f, 0.f, -0.081212f}, br[3] = {0.923122f, 0.924864f, 0.930305f};
    const float ec[3] = {0.846892f, 0.644025f, -0.324398f}, er[3] = {0.276937f, 0.277459f, 0.279092f};
    const float lc[3] = {-ec[0], ec[1], ec[2]};
    MeshSrc m;
    m.subs.push_back(geodesic(5, bc, br));
    m.subs.push_back(geodesic(5, lc, er));
    m.subs.push_back(geodesic(5, ec, er));
    return m;
}

// minimal OBJ reader: `o` starts a submesh, `v` positions, `vt` texture coordinates,
// `f` polygons of v[/vt[/vn]] corners (fan-triangulated)
bool loadObj(const std::string &path, MeshSrc &m, std::string &err)
{
    FILE *fp = fopen(path.c_str(), "r");
    if (!fp) { err = "cannot open " + path; return false; }
    std::vector<float> pos, tex;
    char line[512];
    m.subs.clear();
    while (fgets(line, sizeof(line), fp)) {
        if (line[0] == 'o' && line[1] == ' ') { m.subs.emplace_back(); continue; }
        if (line[0] == 'v' && line[1] == ' ') {
            float x, y, z;
            if (sscanf(line + 2, "%f %f %f", &x, &y, &z) == 3) { pos.push_back(x); pos.push_back(y); pos.push_back(z); }
            continue;
        }
        if (line[0] == 'v' && line[1] == 't' && line[2] == ' ') {
            float u, v;
            if (sscanf(line + 3, "%f %f", &u, &v) == 2) { tex.push_back(u); tex.push_back(v); }
            continue;
        }
        if (line[0] == 'f' && line[1] == ' ') {
            if (m.subs.empty()) m.subs.emplace_back();
            std::vector<int> idx, tidx;
            for (char *p = strtok(line + 2, " \t\r\n"); p; p = strtok(nullptr, " \t\r\n")) {
                int i = atoi(p);
                if (i < 0) i = (int)pos.size() / 3 + i + 1;
                if (i <= 0 || i > (int)pos.size() / 3) { fclose(fp); err = "bad face index in " + path; return false; }
                idx.push_back(i - 1);
                int t = 0;
                if (const char *sl = strchr(p, '/')) t = atoi(sl + 1);
                if (t < 0) t = (int)tex.size() / 2 + t + 1;
                if (t > (int)tex.size() / 2) { fclose(fp); err = "bad texture index in " + path; return false; }
                tidx.push_back(t - 1);
            }
            for (size_t k = 1; k + 1 < idx.size(); k++) {
                Tri t = mkTri(&pos[3 * idx[0]], &pos[3 * idx[k]], &pos[3 * idx[k + 1]]);
                const size_t corner[3] = {0, k, k + 1};
                for (int c = 0; c < 3; c++)
                    if (tidx[corner[c]] >= 0) { t.uv[2 * c] = tex[2 * tidx[corner[c]]]; t.uv[2 * c + 1] = tex[2 * tidx[corner[c]] + 1]; }
                m.subs.back().tris.push_back(t);
            }
        }
    }
    fclose(fp);
    m.subs.erase(std::remove_if(m.subs.begin(), m.subs.end(), [](const SubSrc &s) { return s.tris.empty(); }), m.subs.end());
    if (m.subs.empty()) { err = "no faces in " + path; return false; }
    return true;
}

void boundsOf(const Tri *t, int n, float lo[3], float hi[3])
{
    for (int k = 0; k < 3; k++) { lo[k] = FLT_MAX; hi[k] = -FLT_MAX; }
    for (int i = 0; i < n; i++)
        for (int v = 0; v < 3; v++)
            for (int k = 0; k < 3; k++) {
                lo[k] = std::min(lo[k], t[i].v[3 * v + k]);
                hi[k] = std::max(hi[k], t[i].v[3 * v + k]);
            }
}

// median split on the widest centroid axis until a range holds <= kViewClusterTris triangles
void splitClusters(std::vector<Tri> &tris, int lo, int hi, std::vector<std::pair<int, int>> &out)
{
    if (hi - lo <= kViewClusterTris) { out.push_back({lo, hi - lo}); return; }
    float cmin[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, cmax[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    auto cen = [](const Tri &t, int k) { return t.v[k] + t.v[3 + k] + t.v[6 + k]; };
    for (int i = lo; i < hi; i++)
        for (int k = 0; k < 3; k++) { cmin[k] = std::min(cmin[k], cen(tris[i], k)); cmax[k] = std::max(cmax[k], cen(tris[i], k)); }
    int ax = 0;
    for (int k = 1; k < 3; k++) if (cmax[k] - cmin[k] > cmax[ax] - cmin[ax]) ax = k;
    int mid = (lo + hi) / 2;
    std::nth_element(tris.begin() + lo, tris.begin() + mid, tris.begin() + hi,
                     [&](const Tri &a, const Tri &b) { return cen(a, ax) < cen(b, ax); });
    splitClusters(tris, lo, mid, out);
    splitClusters(tris, mid, hi, out);
}

float srgbToLinear(float s)      // M/include/madrona/render/common.inl:3-10
{
    if (s <= 0.04045f) return s / 12.92f;
    return powf((s + 0.055f) / 1.055f, 2.4f);
}

std::string g_viewErr;

uint32_t be32(const uint8_t *p) { return (uint32_t(p[0]) << 24) | (uint32_t(p[1]) << 16) | (uint32_t(p[2]) << 8) | p[3]; }

}  // namespace

// 8-bit RGB / RGBA, non-interlaced PNG -> RGBA8 rows top to bottom (what the reference's image
// importer hands to cudaMemcpy2DToArray, M/src/render/asset_processor.cpp:312-331).
const char *decodePngRGBA(const char *path, int *width, int *height, std::vector<uint8_t> &rgba)
{
    g_viewErr.clear();
    FILE *fp = fopen(path, "rb");
    if (!fp) { g_viewErr = std::string("cannot open ") + path; return g_viewErr.c_str(); }
    std::vector<uint8_t> file;
    uint8_t buf[65536];
    for (size_t n; (n = fread(buf, 1, sizeof(buf), fp)) > 0;) file.insert(file.end(), buf, buf + n);
    fclose(fp);
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    if (file.size() < 8 || memcmp(file.data(), sig, 8) != 0) { g_viewErr = std::string("not a PNG: ") + path; return g_viewErr.c_str(); }
    int w = 0, h = 0, channels = 0;
    std::vector<uint8_t> idat;
    for (size_t off = 8; off + 12 <= file.size();) {
        uint32_t len = be32(&file[off]);
        const uint8_t *type = &file[off + 4], *data = &file[off + 8];
        if (off + 12 + len > file.size()) break;
        if (!memcmp(type, "IHDR", 4)) {
            w = (int)be32(data); h = (int)be32(data + 4);
            const int depth = data[8], color = data[9], interlace = data[12];
            channels = color == 6 ? 4 : (color == 2 ? 3 : 0);
            if (depth != 8 || channels == 0 || interlace != 0) { g_viewErr = std::string("unsupported PNG format: ") + path; return g_viewErr.c_str(); }
        } else if (!memcmp(type, "IDAT", 4)) {
            idat.insert(idat.end(), data, data + len);
        } else if (!memcmp(type, "IEND", 4)) {
            break;
        }
        off += 12 + len;
    }
    if (w <= 0 || h <= 0 || idat.empty()) { g_viewErr = std::string("truncated PNG: ") + path; return g_viewErr.c_str(); }
    const size_t stride = size_t(w) * channels;
    std::vector<uint8_t> raw((stride + 1) * h);
    uLongf rawLen = (uLongf)raw.size();
    if (uncompress(raw.data(), &rawLen, idat.data(), (uLong)idat.size()) != Z_OK || rawLen != raw.size()) {
        g_viewErr = std::string("PNG inflate failed: ") + path;
        return g_viewErr.c_str();
    }
    // undo the per-row filters (PNG spec 9.2) in place
    std::vector<uint8_t> prev(stride, 0);
    rgba.assign(size_t(w) * h * 4, 255);
    for (int y = 0; y < h; y++) {
        const int filter = raw[(stride + 1) * y];
        uint8_t *row = &raw[(stride + 1) * y + 1];
        for (size_t i = 0; i < stride; i++) {
            const int a = i >= (size_t)channels ? row[i - channels] : 0, b = prev[i];
            const int c = i >= (size_t)channels ? prev[i - channels] : 0;
            int pred = 0;
            switch (filter) {
            case 1: pred = a; break;
            case 2: pred = b; break;
            case 3: pred = (a + b) / 2; break;
            case 4: { int p = a + b - c, pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
                      pred = (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c); } break;
            default: break;
            }
            row[i] = uint8_t(row[i] + pred);
        }
        memcpy(prev.data(), row, stride);
        for (int x = 0; x < w; x++)
            for (int k = 0; k < channels; k++) rgba[(size_t(y) * w + x) * 4 + k] = row[size_t(x) * channels + k];
    }
    *width = w; *height = h;
    return "";
}

const char *buildViewMeshTable(ViewMeshTable &T, ViewTexTable &X, const char *dataDir)
{
    std::memset(&T, 0, sizeof(T));
    std::memset(&X, 0, sizeof(X));
    for (int o = 0; o < OBJ_COUNT; o++) for (int k = 0; k < kViewMaxSubPerObj; k++) X.subTex[o][k] = -1;
    g_viewErr.clear();
    // distinct meshes; Hider and Seeker share one
    enum { M_PLANE, M_CUBE, M_WALL, M_AGENT, M_RAMP, M_BOX, M_COUNT };
    MeshSrc src[M_COUNT];
    if (dataDir && dataDir[0]) {
        const char *names[M_COUNT] = {"plane.obj", "cube_render.obj", "wall_render.obj", "agent_render.obj",
                                      "ramp_render.obj", "elongated_render.obj"};
        for (int i = 0; i < M_COUNT; i++)
            if (!loadObj(std::string(dataDir) + "/" + names[i], src[i], g_viewErr)) return g_viewErr.c_str();
    } else {
        src[M_PLANE] = groundQuad();
        src[M_CUBE] = block(1, 1, -1, 1);
        src[M_WALL] = block(1, 1, 0, 2.5f);
        src[M_AGENT] = agentStandIn();
        src[M_RAMP] = wedge();
        src[M_BOX] = block(4, 0.75f, -1, 1);
    }

    int meshFirstSub[M_COUNT], meshNumSub[M_COUNT];
    float meshAabb[M_COUNT][6];
    // single-cluster meshes first: their triangles are what the kernel stages in shared memory
    auto isSmall = [&](int i) { return src[i].subs.size() == 1 && (int)src[i].subs[0].tris.size() <= kViewSmallTris; };
    std::vector<int> order;
    for (int i = 0; i < M_COUNT; i++) if (isSmall(i)) order.push_back(i);
    const size_t numSmallMeshes = order.size();
    for (int i = 0; i < M_COUNT; i++) if (!isSmall(i)) order.push_back(i);
    for (size_t oi = 0; oi < order.size(); oi++) {
        const int i = order[oi];
        if (oi == numSmallMeshes) T.numSmallTris = T.numTris;
        MeshSrc &m = src[i];
        if ((int)m.subs.size() > kViewMaxSubPerObj || T.numSub + (int)m.subs.size() > kViewMaxSubmeshes) {
            g_viewErr = "render mesh has too many submeshes";
            return g_viewErr.c_str();
        }
        meshFirstSub[i] = T.numSub;
        meshNumSub[i] = (int)m.subs.size();
        for (int k = 0; k < 3; k++) { meshAabb[i][k] = FLT_MAX; meshAabb[i][3 + k] = -FLT_MAX; }
        for (SubSrc &s : m.subs) {
            std::vector<std::pair<int, int>> cl;
            if (isSmall(i)) cl.push_back({0, (int)s.tris.size()});      // one flat list
            else splitClusters(s.tris, 0, (int)s.tris.size(), cl);
            if (cl.size() > 64) { g_viewErr = "render submesh needs more than 64 clusters"; return g_viewErr.c_str(); }
            if (T.numClusters + (int)cl.size() > kViewMaxClusters || T.numTris + (int)s.tris.size() > kViewMaxTris) {
                g_viewErr = "render meshes exceed the view table capacity";
                return g_viewErr.c_str();
            }
            ViewBox &SB = T.sub[T.numSub++];
            SB.first = T.numClusters;
            SB.count = (int)cl.size();
            boundsOf(s.tris.data(), (int)s.tris.size(), SB.lo, SB.hi);
            for (int k = 0; k < 3; k++) {
                meshAabb[i][k] = std::min(meshAabb[i][k], SB.lo[k]);
                meshAabb[i][3 + k] = std::max(meshAabb[i][3 + k], SB.hi[k]);
            }
            for (auto &c : cl) {
                ViewBox &CB = T.cluster[T.numClusters++];
                CB.first = T.numTris + c.first;
                CB.count = c.second;
                boundsOf(s.tris.data() + c.first, c.second, CB.lo, CB.hi);
            }
            for (const Tri &t : s.tris) {
                std::memcpy(X.triUV[T.numTris], t.uv, sizeof(t.uv));
                std::memcpy(T.tri[T.numTris++], t.v, sizeof(t.v));
            }
        }
    }
    if (numSmallMeshes == order.size()) T.numSmallTris = T.numTris;
    // the cluster / submesh boxes cull in object space: pad them like the instance boxes
    auto pad = [](ViewBox &b) {
        for (int k = 0; k < 3; k++) {
            float e = (b.hi[k] - b.lo[k]) * 1e-5f + 1e-6f;
            b.lo[k] -= e; b.hi[k] += e;
        }
    };
    for (int i = 0; i < T.numSub; i++) pad(T.sub[i]);
    for (int i = 0; i < T.numClusters; i++) pad(T.cluster[i]);

    // materials (src/mgr.cpp:624-647); textured ones keep their base colour
    const float grey[3] = {0.4f, 0.4f, 0.4f}, red[3] = {1.0f, 0.1f, 0.1f}, white[3] = {1.f, 1.f, 1.f};
    const float floorc[3] = {0.5f, 0.3f, 0.3f};
    const float ramp[3] = {srgbToLinear(191 / 255.f), srgbToLinear(108 / 255.f), srgbToLinear(10 / 255.f)};
    const float box[3] = {srgbToLinear(12 / 255.f), srgbToLinear(144 / 255.f), srgbToLinear(150 / 255.f)};
    const float eye[3] = {srgbToLinear(230 / 255.f), srgbToLinear(230 / 255.f), srgbToLinear(230 / 255.f)};
    struct Bind { int obj, mesh; const float *col[kViewMaxSubPerObj]; };
    const Bind binds[] = {
        {OBJ_PLANE, M_PLANE, {floorc, floorc, floorc}}, {OBJ_CUBE, M_CUBE, {red, red, red}},
        {OBJ_WALL, M_WALL, {grey, grey, grey}},         {OBJ_HIDER, M_AGENT, {white, eye, eye}},
        {OBJ_SEEKER, M_AGENT, {white, eye, eye}},       {OBJ_RAMP, M_RAMP, {ramp, ramp, ramp}},
        {OBJ_BOX, M_BOX, {box, box, box}},
    };
    for (const Bind &b : binds) {
        T.objFirstSub[b.obj] = meshFirstSub[b.mesh];
        T.objNumSub[b.obj] = meshNumSub[b.mesh];
        std::memcpy(T.objAabb[b.obj], meshAabb[b.mesh], sizeof(float) * 6);
        for (int s = 0; s < kViewMaxSubPerObj; s++)
            for (int k = 0; k < 3; k++) T.subColor[b.obj][s][k] = b.col[s][k];
    }
    // textures (src/mgr.cpp:624-665): image 0 green_grid -> the ground, 1 smile -> hider body,
    // 2 red_smile -> seeker body; only with the reference's data directory (UVs come from its OBJs)
    if (dataDir && dataDir[0]) {
        X.subTex[OBJ_PLANE][0] = 0;
        X.subTex[OBJ_HIDER][0] = 1;
        X.subTex[OBJ_SEEKER][0] = 2;
    }
    // attachEntityToView(..., 100.f, ...) -> fov_scale (M/src/render/ecs_system.cpp:655-660)
    const float toRad = 3.14159265358979323846f / 180.f;
    T.fovScale = 1.0f / tanf(toRad * (100.f * 0.5f));
    return "";
}

}  // namespace hs

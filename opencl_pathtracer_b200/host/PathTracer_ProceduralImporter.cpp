// PathTracer_ProceduralImporter.cpp — see the header for the conventions and the scene list.

#include "PathTracer_ProceduralImporter.h"

#include <cmath>
#include <cstdlib>
#include <functional>
#include <stdexcept>
#include <vector>

namespace PathTracerNS
{
bool Triangle_isValid(Float4 const &s1, Float4 const &s2, Float4 const &s3)
{
    bool ok = (s1.x != s2.x) || (s1.y != s2.y) || (s1.z != s2.z);
    ok &= (s1.x != s3.x) || (s1.y != s3.y) || (s1.z != s3.z);
    ok &= (s2.x != s3.x) || (s2.y != s3.y) || (s2.z != s3.z);
    return ok;
}

// Follows reference Maya/PathTracer_MayaImporter.cpp:943-1047 (tangents/bitangents are not produced: the
// kernel never reads them, Kernel/PathTracer_FullKernel_header.cl:214-215).
void Triangle_Create(Triangle *This, Float4 const &s1, Float4 const &s2, Float4 const &s3, Float2 const &uvp1,
                     Float2 const &uvp2, Float2 const &uvp3, Float2 const &uvn1, Float2 const &uvn2, Float2 const &uvn3,
                     Float4 const &n1, Float4 const &n2, Float4 const &n3, uint positiveMatIndex, uint negativeMatIndex,
                     uint id)
{
    std::memset((void *)This, 0, sizeof(Triangle));
    This->materialWithPositiveNormalIndex = positiveMatIndex;
    This->materialWithNegativeNormalIndex = negativeMatIndex;
    This->id = id;
    BoundingBox_Create(This->AABB, s1, s2, s3);
    const Float4 N = normalize(cross(s2 - s1, s3 - s1));
    This->N = N;

    // order[k] = which input vertex lands in slot k; the six cases of the lexicographic sort
    int order[3];
    if (Vector_LexLessThan(s1, s2))
    {
        if (Vector_LexLessThan(s2, s3)) { order[0] = 0; order[1] = 1; order[2] = 2; }
        else if (Vector_LexLessThan(s1, s3)) { order[0] = 0; order[1] = 2; order[2] = 1; }
        else { order[0] = 2; order[1] = 0; order[2] = 1; }
    }
    else
    {
        if (Vector_LexLessThan(s1, s3)) { order[0] = 1; order[1] = 0; order[2] = 2; }
        else if (Vector_LexLessThan(s2, s3)) { order[0] = 1; order[1] = 2; order[2] = 0; }
        else { order[0] = 2; order[1] = 1; order[2] = 0; }
    }
    const Float4 *S[3] = {&s1, &s2, &s3};
    const Float4 *NV[3] = {&n1, &n2, &n3};
    const Float2 *UP[3] = {&uvp1, &uvp2, &uvp3};
    const Float2 *UN[3] = {&uvn1, &uvn2, &uvn3};
    ptc_float4 *dS[3] = {&This->S1, &This->S2, &This->S3};
    ptc_float4 *dN[3] = {&This->N1, &This->N2, &This->N3};
    ptc_float2 *dUP[3] = {&This->UVP1, &This->UVP2, &This->UVP3};
    ptc_float2 *dUN[3] = {&This->UVN1, &This->UVN2, &This->UVN3};
    for (int k = 0; k < 3; k++)
    {
        *dS[k] = *S[order[k]];
        *dUP[k] = *UP[order[k]];
        *dUN[k] = *UN[order[k]];
        Float4 nv = *NV[order[k]];
        float l = length(nv);
        if (l < 0.5) nv = N;
        else nv = nv / l;
        *dN[k] = nv;
    }
}

namespace
{
const double kPi = 3.14159265358979323846;

inline uint hash32(uint x)
{
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}
inline float hash01(uint a, uint b) { return (float)(hash32(a * 0x9e3779b9u + hash32(b)) >> 8) * (1.0f / 16777216.0f); }

inline Float4 P(double x, double y, double z) { return Float4((float)x, (float)y, (float)z, 1.f); }
inline Float4 V(double x, double y, double z) { return Float4((float)x, (float)y, (float)z, 0.f); }
// Cornell-box data is given in the classic (x right, y up, z depth) frame; the renderer is z-up, and the
// Maya importer maps xyz -> zxy, so the same permutation is used here.
inline Float4 Pc(double x, double y, double z) { return P(z, x, y); }
inline Float4 Vc(double x, double y, double z) { return V(z, x, y); }

struct Vec3d { double x, y, z; };
inline Vec3d crossd(Vec3d a, Vec3d b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline Vec3d subd(Vec3d a, Vec3d b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline double dotd(Vec3d a, Vec3d b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline Vec3d normd(Vec3d a)
{
    double l = std::sqrt(dotd(a, a));
    if (l < 1e-300) return {0, 0, 0};
    return {a.x / l, a.y / l, a.z / l};
}

typedef std::function<Vec3d(double, double)> Surface;

struct SceneBuilder
{
    std::vector<Triangle> tris;
    std::vector<Material> mats;
    std::vector<Light> lights;
    std::vector<Texture> textures;  // scene textures (offsets relative to the start of the scene texels)
    std::vector<Uchar4> texels;     // scene texels
    std::vector<Uchar4> skyTexels;
    Sky sky{};
    Float4 camPos, camDir, camRight, camUp;

    uint addMaterial(int type, float r, float g, float b, float opacity = 1.f, int textureId = -1)
    {
        Material m;
        std::memset((void *)&m, 0, sizeof(m));
        m.type = type;
        m.simpleColor = Float4(r, g, b, 0.f);
        m.textureName = nullptr;
        m.opacity = opacity;
        m.textureId = textureId < 0 ? 0 : textureId;
        m.isSimpleColor = textureId < 0;
        m.hasAlphaMap = false;
        mats.push_back(m);
        return (uint)mats.size() - 1;
    }

    void addTriangle(const Float4 &a, const Float4 &b, const Float4 &c, const Float4 &na, const Float4 &nb,
                     const Float4 &nc, Float2 ua, Float2 ub, Float2 uc, uint mp, uint mn)
    {
        if (!Triangle_isValid(a, b, c)) return;
        // A degenerate vertex normal (pole of a sphere, centre of a disc) would be replaced inside Triangle_Create by the
        // face normal N, whose w lane is 1 (host cross()). The kernel's 4-lane normalize then yields a
        // shading normal with w != 0 and, for axis-aligned faces, a NaN tangent frame (kernel.cl:323) whose NaN ray sweeps
        // the whole BVH in the reference. Scenes are generated without that artefact: such vertices get the face
        // normal as a VECTOR (w = 0), as a modelling package would export it.
        Float4 fn = normalize(cross(b - a, c - a));
        fn.w = 0;
        const Float4 na2 = length(na) < 0.5f ? fn : na, nb2 = length(nb) < 0.5f ? fn : nb, nc2 = length(nc) < 0.5f ? fn : nc;
        Triangle t;
        Triangle_Create(&t, a, b, c, ua, ub, uc, ua, ub, uc, na2, nb2, nc2, mp, mn, (uint)tris.size());
        if (!(t.N.x == t.N.x) || !(t.N.y == t.N.y) || !(t.N.z == t.N.z)) return;  // zero-area sliver
        tris.push_back(t);
    }

    // flat quad a,b,c,d (counter-clockwise seen from the side its normal points to)
    void addQuad(const Float4 &a, const Float4 &b, const Float4 &c, const Float4 &d, uint mp, uint mn)
    {
        Float4 n = normalize(cross(b - a, c - a));
        n.w = 0;
        addTriangle(a, b, c, n, n, n, Float2(0, 0), Float2(1, 0), Float2(1, 1), mp, mn);
        addTriangle(a, c, d, n, n, n, Float2(0, 0), Float2(1, 1), Float2(0, 1), mp, mn);
    }

    // Tessellated parametric surface over [0,1]^2 with smooth vertex normals from central differences.
    // Triangles are wound so that the face normal agrees with du x dv (times -1 if `flip`).
    void addSurface(const Surface &f, int nu, int nv, uint mp, uint mn, bool flip, float uvScale)
    {
        const int W = nu + 1, H = nv + 1;
        std::vector<Float4> pos((size_t)W * H), nor((size_t)W * H);
        const double e = 1e-4;
        for (int j = 0; j < H; j++)
            for (int i = 0; i < W; i++)
            {
                double u = (double)i / nu, v = (double)j / nv;
                Vec3d p = f(u, v);
                Vec3d du = subd(f(u + e, v), f(u - e, v)), dv = subd(f(u, v + e), f(u, v - e));
                Vec3d n = normd(crossd(du, dv));
                if (flip) n = {-n.x, -n.y, -n.z};
                pos[(size_t)j * W + i] = P(p.x, p.y, p.z);
                nor[(size_t)j * W + i] = V(n.x, n.y, n.z);
            }
        for (int j = 0; j < nv; j++)
            for (int i = 0; i < nu; i++)
            {
                size_t a = (size_t)j * W + i, b = a + 1, c = a + W + 1, d = a + W;
                Float2 ua((float)i / nu * uvScale, (float)j / nv * uvScale), ub((float)(i + 1) / nu * uvScale, (float)j / nv * uvScale);
                Float2 uc((float)(i + 1) / nu * uvScale, (float)(j + 1) / nv * uvScale), ud((float)i / nu * uvScale, (float)(j + 1) / nv * uvScale);
                if (!flip)
                {
                    addTriangle(pos[a], pos[b], pos[c], nor[a], nor[b], nor[c], ua, ub, uc, mp, mn);
                    addTriangle(pos[a], pos[c], pos[d], nor[a], nor[c], nor[d], ua, uc, ud, mp, mn);
                }
                else
                {
                    addTriangle(pos[a], pos[c], pos[b], nor[a], nor[c], nor[b], ua, uc, ub, mp, mn);
                    addTriangle(pos[a], pos[d], pos[c], nor[a], nor[d], nor[c], ua, ud, uc, mp, mn);
                }
            }
    }

    void addPointLight(const Float4 &pos, float power, float r, float g, float b)
    {
        Light l;
        std::memset((void *)&l, 0, sizeof(l));
        l.type = LIGHT_POINT;
        l.color = Float4(r, g, b, 0.f);
        l.direction = Float4(0, 0, 0, 1);
        l.position = pos;
        l.power = power;
        lights.push_back(l);
    }
    // PTC_LIGHT_AREA (extension of the CUDA library, include/pathtracer_structs.h ptc_light_make_area: the reference has no area
    // lights): parallelogram corner + a eu + b ev emitting towards cross(eu, ev); ev rides in the lanes a positional light leaves unused
    void addAreaLight(const Float4 &corner, const Float4 &eu, const Float4 &ev, float radiance, float r, float g, float b)
    {
        Light l;
        std::memset((void *)&l, 0, sizeof(l));
        l.type = (LightType)PTC_LIGHT_AREA;
        l.color = Float4(r, g, b, 1.f);
        l.position = Float4(corner.x, corner.y, corner.z, ev.x);
        l.direction = Float4(eu.x, eu.y, eu.z, ev.y);
        l.cosOfInnerFallOffAngle = ev.z;
        l.power = radiance;
        lights.push_back(l);
    }
    void addDirectionalLight(const Float4 &dir, float power, float r, float g, float b)
    {
        Light l;
        std::memset((void *)&l, 0, sizeof(l));
        l.type = LIGHT_DIRECTIONNAL;
        l.color = Float4(r, g, b, 0.f);
        l.direction = normalize(dir);
        l.position = Float4(0, 0, 0, 1);
        l.power = power;
        lights.push_back(l);
    }
    void addSpotLight(const Float4 &pos, const Float4 &dir, float power, double coneAngle, double penumbra, float r, float g, float b)
    {
        Light l;
        std::memset((void *)&l, 0, sizeof(l));
        l.type = LIGHT_SPOT;
        l.color = Float4(r, g, b, 0.f);
        l.cosOfInnerFallOffAngle = (float)std::cos(coneAngle / 2);
        l.cosOfOuterFallOffAngle = (float)std::cos(coneAngle / 2 + penumbra);
        l.direction = normalize(dir);
        l.position = pos;
        l.power = 3 * power;
        lights.push_back(l);
    }

    // pinhole camera: |right| and |up| encode the field of view (reference MayaImporter.cpp:93-100)
    void setCamera(const Float4 &pos, Vec3d dir, Vec3d upHint, double verticalFovDeg, uint w, uint h)
    {
        Vec3d d = normd(dir);
        Vec3d r = normd(crossd(d, upHint));
        Vec3d u = crossd(r, d);
        double upLen = 2.0 * std::tan(verticalFovDeg * kPi / 360.0);
        double rightLen = upLen * (double)w / (double)h;
        camPos = pos;
        camDir = V(d.x, d.y, d.z);
        camRight = V(r.x * rightLen, r.y * rightLen, r.z * rightLen);
        camUp = V(u.x * upLen, u.y * upLen, u.z * upLen);
    }

    uint addTexture(uint w, uint h, const std::function<Uchar4(uint, uint)> &texel)
    {
        Texture t;
        t.width = w;
        t.height = h;
        t.offset = (uint)texels.size();
        texels.reserve(texels.size() + (size_t)w * h);
        for (uint y = 0; y < h; y++)
            for (uint x = 0; x < w; x++) texels.push_back(texel(x, y));
        textures.push_back(t);
        return (uint)textures.size() - 1;
    }

    // six faces of S x S; a vertical gradient with a sun blob, so every face and every texel differs
    void buildSky(bool loadSky, uint S)
    {
        sky.cosRotationAngle = 1;
        sky.sinRotationAngle = 0;
        sky.exposantFactorX = 0;
        sky.exposantFactorY = 0;
        sky.groundScale = 1;
        if (!loadSky)
        {
            for (int i = 0; i < 6; i++) { sky.skyTextures[i].width = 1; sky.skyTextures[i].height = 1; sky.skyTextures[i].offset = 0; }
            skyTexels.assign(1, Uchar4(0, 0, 0, 0));
            return;
        }
        skyTexels.clear();
        for (int f = 0; f < 6; f++)
        {
            sky.skyTextures[f].width = S;
            sky.skyTextures[f].height = S;
            sky.skyTextures[f].offset = (uint)skyTexels.size();
            for (uint y = 0; y < S; y++)
                for (uint x = 0; x < S; x++)
                {
                    float u = (float)x / (S - 1), v = (float)y / (S - 1);
                    float top = f == 5 ? 1.f : (f == 0 ? 0.f : v);
                    float rr = 0.35f + 0.25f * top + 0.1f * u, gg = 0.5f + 0.3f * top, bb = 0.65f + 0.35f * top;
                    if (f == 0) { rr = 0.25f; gg = 0.22f; bb = 0.2f; }
                    if (f == 5) { float dx = u - 0.6f, dy = v - 0.4f; if (dx * dx + dy * dy < 0.01f) { rr = gg = bb = 1.f; } }
                    skyTexels.push_back(Uchar4((uchar)(255 * rr > 255 ? 255 : 255 * rr), (uchar)(255 * gg > 255 ? 255 : 255 * gg),
                                               (uchar)(255 * bb > 255 ? 255 : 255 * bb), (uchar)((x + y + f) & 1 ? 255 : 0)));
                }
        }
    }
};

// ------------------------------------------------------------------------------------------------------
// Cornell box (classic measured data, frame permuted to z-up)
void buildCornell(SceneBuilder &s, bool specular, uint w, uint h, bool areaLight = false)
{
    uint white = s.addMaterial(MAT_STANDART, 0.73f, 0.73f, 0.73f);
    uint red = s.addMaterial(MAT_STANDART, 0.65f, 0.05f, 0.05f);
    uint green = s.addMaterial(MAT_STANDART, 0.12f, 0.45f, 0.15f);
    uint tall = white, shortm = white;
    if (specular)
    {
        tall = s.addMaterial(MAT_GLASS, 0.9f, 0.95f, 1.0f, 0.1f);
        shortm = s.addMaterial(MAT_VARNHISHED, 0.3f, 0.35f, 0.7f);
    }
    // floor, ceiling, back wall, right (green), left (red); normals point into the room
    s.addQuad(Pc(552.8, 0, 0), Pc(0, 0, 0), Pc(0, 0, 559.2), Pc(549.6, 0, 559.2), white, white);
    s.addQuad(Pc(556.0, 548.8, 0), Pc(556.0, 548.8, 559.2), Pc(0, 548.8, 559.2), Pc(0, 548.8, 0), white, white);
    s.addQuad(Pc(549.6, 0, 559.2), Pc(0, 0, 559.2), Pc(0, 548.8, 559.2), Pc(556.0, 548.8, 559.2), white, white);
    s.addQuad(Pc(0, 0, 559.2), Pc(0, 0, 0), Pc(0, 548.8, 0), Pc(0, 548.8, 559.2), green, green);
    s.addQuad(Pc(552.8, 0, 0), Pc(549.6, 0, 559.2), Pc(556.0, 548.8, 559.2), Pc(556.0, 548.8, 0), red, red);
    // short block
    const double sb[4][2] = {{130, 65}, {82, 225}, {240, 272}, {290, 114}};
    s.addQuad(Pc(sb[0][0], 165, sb[0][1]), Pc(sb[1][0], 165, sb[1][1]), Pc(sb[2][0], 165, sb[2][1]), Pc(sb[3][0], 165, sb[3][1]), shortm, shortm);
    for (int i = 0; i < 4; i++)
    {
        int j = (i + 3) % 4;
        s.addQuad(Pc(sb[i][0], 0, sb[i][1]), Pc(sb[i][0], 165, sb[i][1]), Pc(sb[j][0], 165, sb[j][1]), Pc(sb[j][0], 0, sb[j][1]), shortm, shortm);
    }
    // tall block
    const double tb[4][2] = {{423, 247}, {265, 296}, {314, 456}, {472, 406}};
    s.addQuad(Pc(tb[0][0], 330, tb[0][1]), Pc(tb[1][0], 330, tb[1][1]), Pc(tb[2][0], 330, tb[2][1]), Pc(tb[3][0], 330, tb[3][1]), tall, tall);
    for (int i = 0; i < 4; i++)
    {
        int j = (i + 3) % 4;
        s.addQuad(Pc(tb[i][0], 0, tb[i][1]), Pc(tb[i][0], 330, tb[i][1]), Pc(tb[j][0], 330, tb[j][1]), Pc(tb[j][0], 0, tb[j][1]), tall, tall);
    }
    // the reference has no area light: a point light just under the ceiling
    // "cornell_area" (BASELINE.json configs[1] names an area light): the classic 130 x 105 emitter, a hair under the ceiling
    if (areaLight) s.addAreaLight(Pc(213, 548.7, 227), Vc(130, 0, 0), Vc(0, 0, 105), 100.f, 1.f, 0.96f, 0.9f);
    else s.addPointLight(Pc(278, 540, 279.5), 250000.f, 1.f, 0.96f, 0.9f);
    // classic camera: (278,273,-800) looking +z, 35 mm lens on a 25 mm film; screen-right is -x
    Float4 pos = Pc(278, 273, -800);
    Float4 d = Vc(0, 0, 1), up = Vc(0, 1, 0);
    s.setCamera(pos, {d.x, d.y, d.z}, {up.x, up.y, up.z}, 2.0 * std::atan(0.0125 / 0.035) * 180.0 / kPi, w, h);
}

// ------------------------------------------------------------------------------------------------------
// "gallery": small scene touching every kernel branch
void buildGallery(SceneBuilder &s, uint seed, uint w, uint h)
{
    uint tex0 = s.addTexture(64, 64, [&](uint x, uint y) {
        bool c = ((x / 8) + (y / 8)) & 1;
        return Uchar4(c ? 230 : 40, c ? 220 : 60, c ? 200 : 90, (uchar)(hash32(x * 64 + y + seed) & 255));
    });
    uint tex1 = s.addTexture(32, 48, [&](uint x, uint y) { return Uchar4((uchar)(x * 8), (uchar)(y * 5), (uchar)(255 - x * 4), 0); });
    uint stone = s.addMaterial(MAT_STANDART, 0.7f, 0.68f, 0.6f);
    uint checker = s.addMaterial(MAT_STANDART, 1, 1, 1, 1.f, (int)tex0);
    uint grad = s.addMaterial(MAT_VARNHISHED, 1, 1, 1, 1.f, (int)tex1);
    uint glass = s.addMaterial(MAT_GLASS, 0.85f, 0.95f, 0.9f, 0.15f);
    uint water = s.addMaterial(MAT_WATER, 0.8f, 0.9f, 1.0f);
    uint varnish = s.addMaterial(MAT_VARNHISHED, 0.75f, 0.2f, 0.15f);
    uint metal = s.addMaterial(MAT_METAL, 0.9f, 0.9f, 0.9f);
    // pool floor (checker), three walls, water surface, objects
    s.addSurface([](double u, double v) { return Vec3d{u * 400, v * 400, 6 * std::sin(u * 9) * std::cos(v * 7)}; }, 12, 12, checker, stone, false, 4.f);
    s.addSurface([](double u, double v) { return Vec3d{u * 400, 400 + 4 * std::sin(u * 12 + v * 5), v * 250}; }, 10, 8, grad, stone, true, 1.f);
    s.addSurface([](double u, double v) { return Vec3d{0, u * 400, v * 250}; }, 6, 6, stone, stone, false, 1.f);
    s.addSurface([](double u, double v) { return Vec3d{400, u * 400, v * 250}; }, 6, 6, varnish, stone, true, 1.f);
    s.addSurface([](double u, double v) { return Vec3d{40 + u * 320, 40 + v * 320, 30 + 3 * std::sin(u * 20) * std::sin(v * 17)}; }, 14, 14, water, water, false, 1.f);
    auto sphere = [](double cx, double cy, double cz, double r) {
        return [=](double u, double v) {
            double th = u * 2 * kPi, ph = (v - 0.5) * kPi;
            return Vec3d{cx + r * std::cos(ph) * std::cos(th), cy + r * std::cos(ph) * std::sin(th), cz + r * std::sin(ph)};
        };
    };
    s.addSurface(sphere(130, 150, 90, 50), 16, 10, glass, glass, false, 1.f);
    s.addSurface(sphere(270, 230, 80, 45), 16, 10, varnish, varnish, false, 1.f);
    s.addSurface(sphere(200, 320, 70, 35), 12, 8, metal, metal, false, 1.f);
    s.addSurface(sphere(300, 110, 60, 30), 12, 8, checker, checker, false, 2.f);
    s.addDirectionalLight(V(0.35, 0.45, -0.82), 1.6f, 1.f, 0.95f, 0.85f);
    s.addPointLight(P(200, 200, 230), 60000.f, 0.9f, 0.9f, 1.f);
    s.addSpotLight(P(60, 60, 240), V(0.5, 0.55, -0.67), 0.9f, 0.7, 0.25, 1.f, 0.8f, 0.6f);
    s.setCamera(P(200, -330, 190), {0, 1, -0.28}, {0, 0, 1}, 50.0, w, h);
}

// ------------------------------------------------------------------------------------------------------
// "atrium": Sponza-scale hall. k divides the tessellation, m multiplies it (k = m = 1 -> ~262 k triangles; "atrium:4" -> ~16 k;
// "atrium@2" -> ~1.05 M, the north star's target size: same surfaces, materials and lights, every grid twice as fine).
void buildAtrium(SceneBuilder &s, uint seed, int k, int m, uint w, uint h)
{
    auto R = [&](int n) { int v = n * m / k; return v < 2 ? 2 : v; };
    const double LX = 3000, LY = 1800, LZ = 1200;
    uint wall = s.addMaterial(MAT_STANDART, 0.75f, 0.70f, 0.62f);
    uint floorM = s.addMaterial(MAT_VARNHISHED, 0.60f, 0.55f, 0.50f);
    uint column = s.addMaterial(MAT_STANDART, 0.80f, 0.78f, 0.72f);
    uint glass = s.addMaterial(MAT_GLASS, 0.90f, 0.95f, 1.00f, 0.1f);
    uint drapeA = s.addMaterial(MAT_STANDART, 0.70f, 0.12f, 0.10f);
    uint drapeB = s.addMaterial(MAT_VARNHISHED, 0.15f, 0.45f, 0.25f);
    uint ceilM = s.addMaterial(MAT_STANDART, 0.82f, 0.82f, 0.80f);
    float ph[8];
    for (int i = 0; i < 8; i++) ph[i] = 6.2831853f * hash01(seed, (uint)i);

    // floor (varnished stone, gentle bumps) — normal +z
    s.addSurface([&](double u, double v) {
        return Vec3d{u * LX, v * LY, 2.5 * std::sin(u * 97 + ph[0]) * std::cos(v * 61 + ph[1]) + 1.5 * std::sin(u * 211 + v * 173 + ph[2])};
    }, R(160), R(96), floorM, floorM, false, 8.f);
    // two aisle ceilings (the nave is open to the sky) — normal -z
    for (int side = 0; side < 2; side++)
    {
        double y0 = side == 0 ? 0 : LY - 500;
        s.addSurface([&, y0](double u, double v) {
            return Vec3d{u * LX, y0 + v * 500, LZ - 6 * std::sin(u * 60 + ph[3]) * std::sin(v * 9)};
        }, R(160), R(28), ceilM, ceilM, true, 4.f);
    }
    // side walls y = 0 (normal +y) and y = LY (normal -y), brick relief
    s.addSurface([&](double u, double v) {
        return Vec3d{u * LX, 4 * std::sin(u * 260 + ph[4]) * std::sin(v * 120), v * LZ};
    }, R(160), R(64), wall, wall, true, 6.f);
    s.addSurface([&](double u, double v) {
        return Vec3d{u * LX, LY - 4 * std::sin(u * 260 + ph[5]) * std::sin(v * 120), v * LZ};
    }, R(160), R(64), wall, wall, false, 6.f);
    // end walls x = 0 (normal +x) and x = LX (normal -x)
    s.addSurface([&](double u, double v) {
        return Vec3d{4 * std::sin(u * 150 + ph[6]) * std::sin(v * 120), u * LY, v * LZ};
    }, R(96), R(64), wall, wall, false, 6.f);
    s.addSurface([&](double u, double v) {
        return Vec3d{LX - 4 * std::sin(u * 150 + ph[7]) * std::sin(v * 120), u * LY, v * LZ};
    }, R(96), R(64), wall, wall, true, 6.f);
    // two rows of 8 fluted columns; four of them glass
    for (int row = 0; row < 2; row++)
        for (int c = 0; c < 8; c++)
        {
            double cx = 300 + 340.0 * c, cy = row == 0 ? 500 : LY - 500;
            bool isGlass = (row == 0 && (c == 2 || c == 5)) || (row == 1 && (c == 1 || c == 6));
            uint m = isGlass ? glass : column;
            s.addSurface([=](double u, double v) {
                double th = u * 2 * kPi;
                double r = 45 + 4 * std::cos(16 * th) + 10 * std::exp(-v * 30) + 10 * std::exp(-(1 - v) * 30);
                return Vec3d{cx + r * std::cos(th), cy + r * std::sin(th), v * 900};
            }, R(48), R(40), m, m, false, 1.f);
            // cap
            s.addSurface([=](double u, double v) {
                double th = u * 2 * kPi, r = 59 * (1 - v);
                return Vec3d{cx + r * std::cos(th), cy + r * std::sin(th), 900};
            }, R(48), 1, m, m, false, 1.f);
        }
    // arches between consecutive columns (half tori)
    for (int row = 0; row < 2; row++)
        for (int c = 0; c < 7; c++)
        {
            double cx = 300 + 340.0 * c + 170, cy = row == 0 ? 500 : LY - 500;
            s.addSurface([=](double u, double v) {
                double a = u * kPi, b = v * 2 * kPi;
                double rr = 170 + 35 * std::cos(b);
                return Vec3d{cx - rr * std::cos(a), cy + 35 * std::sin(b), 900 + rr * std::sin(a)};
            }, R(32), R(24), column, column, true, 1.f);
        }
    // drapes across the nave
    for (int d = 0; d < 8; d++)
    {
        double x0 = 380 + 320.0 * d;
        uint m = (d & 1) ? drapeB : drapeA;
        float p0 = 6.2831853f * hash01(seed + 17, (uint)d);
        s.addSurface([=](double u, double v) {
            double sag = 160 * 4 * v * (1 - v);
            return Vec3d{x0 + u * 180 + 12 * std::sin(v * 25 + p0), 520 + v * (LY - 1040), 1020 - sag + 14 * std::sin(u * 19 + v * 31 + p0)};
        }, R(64), R(48), m, m, false, 3.f);
    }
    // four glass blobs on the floor
    for (int b = 0; b < 4; b++)
    {
        double cx = 700 + 520.0 * b, cy = 900 + ((b & 1) ? 140 : -140);
        float p0 = 6.2831853f * hash01(seed + 31, (uint)b);
        s.addSurface([=](double u, double v) {
            double th = u * 2 * kPi, phi = (v - 0.5) * kPi;
            double r = 70 + 9 * std::sin(5 * th + p0) * std::cos(3 * phi) + 6 * std::sin(9 * phi + p0);
            return Vec3d{cx + r * std::cos(phi) * std::cos(th), cy + r * std::cos(phi) * std::sin(th), 78 + r * std::sin(phi)};
        }, R(48), R(40), glass, glass, false, 1.f);
    }
    s.addPointLight(P(900, 900, 1000), 1.6e6f, 1.f, 0.93f, 0.85f);
    s.addPointLight(P(2100, 900, 1000), 1.6e6f, 0.85f, 0.9f, 1.f);
    s.addDirectionalLight(V(0.30, 0.25, -0.92), 2.0f, 1.f, 0.97f, 0.9f);
    s.setCamera(P(200, 900, 170), {1, 0.05, 0.12}, {0, 0, 1}, 45.0, w, h);
}

// ------------------------------------------------------------------------------------------------------
// "terrain": g x g flattened instances of a displaced rock over a displaced terrain, textured, sun + sky.
void buildTerrain(SceneBuilder &s, uint seed, int g, uint w, uint h)
{
    const double L = 40.0 * 64;  // world side, independent of g so that distances stay >> 1
    const int terrainRes = (int)(1000.0 * g / 64.0) < 8 ? 8 : (int)(1000.0 * g / 64.0);
    const uint texRes = g >= 64 ? 1024u : (g >= 16 ? 256u : 64u);
    float ph[6];
    for (int i = 0; i < 6; i++) ph[i] = 6.2831853f * hash01(seed + 99, (uint)i);
    std::vector<uint> mats;
    for (uint t = 0; t < 16; t++)
    {
        uint id = s.addTexture(texRes, texRes, [&](uint x, uint y) {
            uint hsh = hash32((x / 4) * 7919u + (y / 4) * 104729u + t * 31u + seed);
            float base = 0.35f + 0.4f * (float)((hsh >> 8) & 255) / 255.f;
            float rr = base * (0.6f + 0.4f * (float)((t * 37) % 16) / 15.f), gg = base * (0.5f + 0.5f * (float)((t * 11) % 16) / 15.f),
                  bb = base * (0.4f + 0.5f * (float)((t * 5) % 16) / 15.f);
            return Uchar4((uchar)(rr * 255), (uchar)(gg * 255), (uchar)(bb * 255), 0);
        });
        mats.push_back(s.addMaterial(t % 4 == 3 ? MAT_VARNHISHED : MAT_STANDART, 1, 1, 1, 1.f, (int)id));
    }
    auto height = [=](double x, double y) {
        return 60 * std::sin(x * 0.004 + ph[0]) * std::cos(y * 0.0035 + ph[1]) + 25 * std::sin(x * 0.013 + y * 0.011 + ph[2]) +
               6 * std::sin(x * 0.05 + ph[3]) * std::sin(y * 0.047 + ph[4]);
    };
    s.addSurface([=](double u, double v) { return Vec3d{u * L, v * L, height(u * L, v * L)}; }, terrainRes, terrainRes, mats[0], mats[0], false, 64.f);
    const double cell = L / g;
    for (int j = 0; j < g; j++)
        for (int i = 0; i < g; i++)
        {
            uint inst = (uint)(j * g + i);
            double cx = (i + 0.5) * cell + (hash01(seed + 1, inst) - 0.5) * cell * 0.4;
            double cy = (j + 0.5) * cell + (hash01(seed + 2, inst) - 0.5) * cell * 0.4;
            double r0 = cell * (0.16 + 0.1 * hash01(seed + 3, inst));
            double cz = height(cx, cy) + r0 * 0.6;
            float p0 = 6.2831853f * hash01(seed + 4, inst);
            uint m = mats[1 + inst % 15];
            s.addSurface([=](double u, double v) {
                double th = u * 2 * kPi, phi = (v - 0.5) * kPi;
                double r = r0 * (1 + 0.18 * std::sin(4 * th + p0) * std::cos(3 * phi) + 0.1 * std::sin(7 * phi + 2 * th + p0));
                return Vec3d{cx + r * std::cos(phi) * std::cos(th), cy + r * std::cos(phi) * std::sin(th), cz + r * std::sin(phi)};
            }, 31, 32, m, m, false, 2.f);
        }
    s.addDirectionalLight(V(0.45, 0.30, -0.84), 2.2f, 1.f, 0.96f, 0.88f);
    s.setCamera(P(L * 0.5, -L * 0.08, 260), {0.0, 1.0, -0.22}, {0, 0, 1}, 50.0, w, h);
}

}  // namespace

void PathTracerProceduralImporter::Import(uint image_width, uint image_height, bool loadSky)
{
    *ptr__global__imageWidth = image_width;
    *ptr__global__imageHeight = image_height;
    *ptr__global__imageSize = image_width * image_height;

    std::string name = sceneName;
    int param = 0;
    int refine = 1;  // "name@m": finer tessellation (atrium only)
    size_t at = name.find('@');
    if (at != std::string::npos)
    {
        refine = std::atoi(name.c_str() + at + 1);
        if (refine < 1 || refine > 8) throw std::runtime_error("PathTracerProceduralImporter: refinement must be 1..8 in '" + sceneName + "'");
        name = name.substr(0, at);
    }
    size_t colon = name.find(':');
    if (colon != std::string::npos)
    {
        param = std::atoi(name.c_str() + colon + 1);
        name = name.substr(0, colon);
    }

    SceneBuilder s;
    uint skySize = 1;
    if (name == "cornell") buildCornell(s, false, image_width, image_height);
    else if (name == "cornell_specular") buildCornell(s, true, image_width, image_height);
    else if (name == "cornell_area") buildCornell(s, false, image_width, image_height, true);
    else if (name == "gallery") { buildGallery(s, seed, image_width, image_height); skySize = 32; }
    else if (name == "atrium") buildAtrium(s, seed, param > 0 ? param : 1, refine, image_width, image_height);
    else if (name == "terrain") { int g = param > 0 ? param : 64; buildTerrain(s, seed, g, image_width, image_height); skySize = g >= 64 ? 512 : 64; }
    else throw std::runtime_error("PathTracerProceduralImporter: unknown scene '" + sceneName + "'");
    s.buildSky(loadSky, skySize);

    if (s.tris.empty()) throw std::runtime_error("Scene is empty. Stopping.");  // as reference MayaImporter.cpp:39-40

    *ptr__global__cameraPosition = s.camPos;
    *ptr__global__cameraDirection = s.camDir;
    *ptr__global__cameraRight = s.camRight;
    *ptr__global__cameraUp = s.camUp;

    const size_t nt = s.tris.size();
    *ptr__global__triangulation = new Triangle[nt];
    std::memcpy((void *)*ptr__global__triangulation, (void *)s.tris.data(), nt * sizeof(Triangle));
    *ptr__global__triangulationSize = (uint)nt;

    *ptr__global__lights = new Light[s.lights.size() + 1];
    if (!s.lights.empty()) std::memcpy((void *)*ptr__global__lights, (void *)s.lights.data(), s.lights.size() * sizeof(Light));
    *ptr__global__lightsSize = (uint)s.lights.size();

    *ptr__global__materiaux = new Material[s.mats.size() + 1];
    std::memcpy((void *)*ptr__global__materiaux, (void *)s.mats.data(), s.mats.size() * sizeof(Material));
    *ptr__global__materiauxSize = (uint)s.mats.size();

    // texels: sky faces first, scene textures after them (reference MayaImporter LoadSkyAndAllocateTextureMemory)
    const uint skyCount = (uint)s.skyTexels.size();
    *ptr__global__textures = new Texture[s.textures.size() + 1];
    for (size_t i = 0; i < s.textures.size(); i++)
    {
        (*ptr__global__textures)[i] = s.textures[i];
        (*ptr__global__textures)[i].offset += skyCount;
    }
    *ptr__global__texturesSize = (uint)s.textures.size();
    *ptr__global__texturesDataSize = skyCount + (uint)s.texels.size();
    *ptr__global__texturesData = new Uchar4[*ptr__global__texturesDataSize];
    std::memcpy((void *)*ptr__global__texturesData, (void *)s.skyTexels.data(), skyCount * sizeof(Uchar4));
    if (!s.texels.empty())
        std::memcpy((void *)(*ptr__global__texturesData + skyCount), (void *)s.texels.data(), s.texels.size() * sizeof(Uchar4));
    *ptr__global__sky = s.sky;
}

}  // namespace PathTracerNS

// ORACLE — test infrastructure only (see oracle_math.h header).  Parity status: pinned to the shipped SPIR-V (tests/golden/spirv_frames.npz); driver-defined parts unpinned.
//
// CPU restatement of the live shader path of ARTurleite6/RayTracer, one function per shader
// function, each citing the file:line it follows:
//   shaders/simple.rgen   (ray generation, path loop, accumulation)
//   shaders/simple.rchit  (hit shading: NEE + MIS, GGX/Lambert BSDF sampling, emissive rules)
//   shaders/simple.rmiss  (miss)
//   shaders/restir.rgen / restir.rchit / restir.rmiss (G-buffer pass)
// GLSL built-ins with driver-defined precision are replaced by the arithmetic contract in
// oracle_math.h.  Behaviour the reference leaves undefined is DEFINED here:
//   * weights[i] of the skipped light is 0 (simple.rchit:555-574 leaves it uninitialised)
//   * TerminateOnFirstHit visibility = "any hit in range whose instance is not the light"
//   * lights.length() == 0 when there are no lights (the reference fails to create the buffer)
#include <cstdio>
#include <vector>

#include "oracle_scene.h"

namespace orc {

struct Cfg {
    int max_depth;        // simple.rgen:23 MAX_DEPTH = 3
    float clamp;          // simple.rgen:121 min(color, 5.0)
    int use_bvh;          // 1 = BVH8 traversal, 0 = brute force
};

static const int MAXLIGHTS = 256;        // simple.rchit:13
static const float EPS_PDF = 1e-6f;      // simple.rchit:63
static const float EPS_COS = 1e-4f;      // simple.rchit:64
static const float EPS_VOH = 1e-4f;      // simple.rchit:65
static const float MIN_ROUGHNESS = 0.02f;  // simple.rchit:66

struct Payload {  // ray_common.glsl:13-26
    V3 color, throughput, hitPosition, nextDirection;
    uint32_t seed;
    bool hit, firstBounce, isSpecular;
    float previousBrdfPdf;
    V3 previousHitPos;
    float p_sample_light;
    bool didDirectIllumination;
};

struct Basis { V3 b0, b1, b2; };

// math.glsl:9-15
static inline Basis createBasis(V3 normal) {
    V3 axis2 = normalize(normal);
    V3 a = fabsf(axis2.x) > 0.9f ? v3(0, 1, 0) : v3(1, 0, 0);
    V3 axis1 = normalize(cross(axis2, a));
    V3 axis0 = cross(axis2, axis1);
    return Basis{axis0, axis1, axis2};
}
// math.glsl:18-24
static inline V3 worldToLocal(V3 v, const Basis& b) { return v3(dot(v, b.b0), dot(v, b.b1), dot(v, b.b2)); }
// math.glsl:27-29
static inline V3 localToWorld(V3 v, const Basis& b) { return (b.b0 * v.x + b.b1 * v.y) + b.b2 * v.z; }
static inline float max3(V3 v) { return gmax(v.x, gmax(v.y, v.z)); }  // math.glsl:39-41

static inline V3 albedo_of(const Material& m) { return v3(m.albedo[0], m.albedo[1], m.albedo[2]); }
static inline V3 F0_of(const Material& m) { return mix3(v3s(0.04f), albedo_of(m), m.metallic); }

// simple.rchit:69-75
static inline float getSpecularProbability(const Material& m) {
    float baseProb = max3(F0_of(m));
    float ri = smoothstep(0.0f, 1.0f, m.roughness * 0.7f);
    return mixf(baseProb, baseProb * 0.5f, ri);
}
// simple.rchit:77-83
static inline float D_GGX(float NoH, float roughness) {
    float a = gmax(roughness, MIN_ROUGHNESS);
    float a2 = a * a;
    float nh = gclamp(NoH, 0.0f, 1.0f);
    float denom = (nh * nh) * (a2 - 1.0f) + 1.0f;
    return a2 / ((kPi * denom) * denom);
}
// simple.rchit:85-93
static inline float G_Smith(float NoV, float NoL, float roughness) {
    float a = gmax(roughness, MIN_ROUGHNESS);
    float k = a * 0.5f;
    float nv = gclamp(NoV, EPS_COS, 1.0f);
    float nl = gclamp(NoL, EPS_COS, 1.0f);
    float G1V = nv / (nv * (1.0f - k) + k);
    float G1L = nl / (nl * (1.0f - k) + k);
    return G1V * G1L;
}
// simple.rchit:96-98
static inline V3 F_Schlick(V3 F0, float VoH) { return F0 + (v3s(1.0f) - F0) * pow5(1.0f - VoH); }
// simple.rchit:100-110
static inline V3 generateCosineWeightedDirection(float rx, float ry) {
    float phi = kTwoPi * rx;
    float cosTheta = sqrtf(ry);
    float sinTheta = sqrtf(1.0f - ry);
    float s, c;
    sincos_contract(phi, &s, &c);
    return v3(c * sinTheta, s * sinTheta, cosTheta);
}
// diffuse lobe value used by sampleBRDF (simple.rchit:416-422, 434-440)
static inline V3 diffuseValue(V3 wo, V3 wi, const Material& m) {
    V3 h = normalize(wo + wi);
    float VoH = gclamp(dot(wo, h), 0.0f, 1.0f);
    V3 F = F_Schlick(F0_of(m), VoH);
    V3 diffuseAlbedo = albedo_of(m) * (1.0f - m.metallic);
    return (diffuseAlbedo * (v3s(1.0f) - F)) / kPi;
}
// simple.rchit:118-166 (evaluateBRDFComponents + evaluateFullBRDF; the pdf members are unused)
static inline V3 evaluateFullBRDF(V3 wo, V3 wi, const Material& m) {
    float NoL = wi.z, NoV = wo.z;
    if (NoL <= 0.0f || NoV <= 0.0f) return v3s(0.0f);
    V3 h = normalize(wo + wi);
    float NoH = h.z;
    float VoH = dot(wo, h);
    V3 F = F_Schlick(F0_of(m), VoH);
    float nonMetalWeight = 1.0f - m.metallic;
    V3 diffuseAlbedo = albedo_of(m) * nonMetalWeight;
    V3 diffuse = (diffuseAlbedo * (v3s(1.0f) - F)) / kPi;
    float D = D_GGX(NoH, m.roughness);
    float G = G_Smith(NoV, NoL, m.roughness);
    V3 specular = (F * (D * G)) / ((4.0f * NoV) * NoL);
    return diffuse + specular;
}
// simple.rchit:168-193
static inline V3 microfacetF(V3 wo, V3 wi, V3 h, const Material& m) {
    float NoL = wi.z, NoV = wo.z;
    if (NoL <= 0.0f || NoV <= 0.0f) return v3s(0.0f);
    float NoH = h.z;
    float VoH = dot(wo, h);
    float D = D_GGX(NoH, m.roughness);
    float G = G_Smith(NoV, NoL, m.roughness);
    V3 F = F_Schlick(F0_of(m), VoH);
    return (F * (D * G)) / ((4.0f * NoV) * NoL);
}
// simple.rchit:195-200
static inline float microfacetPDF(V3 wo, V3 h, float roughness) {
    float nh = gmax(h.z, EPS_COS);
    float voh = gmax(dot(wo, h), EPS_VOH);
    float D = D_GGX(nh, roughness);
    return gmax((D * nh) / (4.0f * voh), EPS_PDF);
}
// simple.rchit:202-217
static inline V3 sampleGGX(float rx, float ry, float roughness) {
    float a = roughness * roughness;
    float phi = kTwoPi * rx;
    float cosTheta = sqrtf((1.0f - ry) / (1.0f + (a * a - 1.0f) * ry));
    float sinTheta = sqrtf(1.0f - cosTheta * cosTheta);
    float s, c;
    sincos_contract(phi, &s, &c);
    return v3(sinTheta * c, sinTheta * s, cosTheta);
}
// simple.rchit:234-237
static inline float misWeightPower(float pdf1, float pdf2) {
    if (pdf1 <= 0.0f || pdf2 <= 0.0f) return 0.0f;
    return (pdf1 * pdf1) / (pdf1 * pdf1 + pdf2 * pdf2);
}

struct Ctx {
    const Scene* s;
    Cfg cfg;
    Counters* cnt;
};

static inline V3 vpos(const Vertex& v) { return v3(v.pos[0], v.pos[1], v.pos[2]); }
static inline V3 vnrm(const Vertex& v) { return v3(v.normal[0], v.normal[1], v.normal[2]); }

struct LightSample {
    V3 position, normal, emission, direction;
    float pdf, distance;
    bool valid;
};

// simple.rchit:239-322
static LightSample sampleLight(const Ctx& cx, uint32_t lightIdx, V3 hitPos, uint32_t& seed) {
    LightSample r; r.valid = false;
    const Scene& s = *cx.s;
    if (lightIdx >= s.lights.size()) return r;
    const Light& light = s.lights[lightIdx];
    const Object& lo = s.objects[light.object_index];
    const Material& lm = s.materials[lo.material_index];
    const Mesh& mesh = s.meshes[lo.mesh_index];
    uint32_t numTriangles = light.num_triangles;
    uint32_t triangleIdx = (uint32_t)(rnd(seed) * (float)numTriangles);
    if (numTriangles == 0u) {  // DEFINED: a light whose mesh is empty (the shader would index triangle 0xFFFFFFFF) gives an
        rnd(seed); rnd(seed);  // invalid sample; the two point draws are still taken
        return r;
    }
    if (triangleIdx > numTriangles - 1u) triangleIdx = numTriangles - 1u;
    const Vertex& v0 = mesh.v[mesh.idx[3 * triangleIdx]];
    const Vertex& v1 = mesh.v[mesh.idx[3 * triangleIdx + 1]];
    const Vertex& v2 = mesh.v[mesh.idx[3 * triangleIdx + 2]];
    float r1 = rnd(seed);
    float r2 = rnd(seed);
    float sqrtR1 = sqrtf(r1);
    float u = 1.0f - sqrtR1;
    float v = sqrtR1 * (1.0f - r2);
    float w = sqrtR1 * r2;
    V3 localPos = (vpos(v0) * u + vpos(v1) * v) + vpos(v2) * w;
    V3 worldV0 = xform_point(light.transform, vpos(v0));
    V3 worldV1 = xform_point(light.transform, vpos(v1));
    V3 worldV2 = xform_point(light.transform, vpos(v2));
    V3 worldPos = xform_point(light.transform, localPos);
    V3 crossE = cross(worldV1 - worldV0, worldV2 - worldV0);
    V3 worldNormal = normalize(crossE);
    V3 toSurface = hitPos - worldPos;
    float cosL = dot(worldNormal, normalize(toSurface));
    if (cosL < 0.0f) { cosL = fabsf(cosL); worldNormal = neg(worldNormal); }
    if (cosL > 0.0f) {
        r.position = worldPos;
        r.normal = worldNormal;
        V3 toLight = r.position - hitPos;
        float safeDist = gmax(length(toLight), 0.01f);
        r.distance = safeDist;
        r.direction = toLight / r.distance;
        float triangleArea = 0.5f * length(crossE);
        float cosTheta = gmax(0.0f, dot(neg(r.direction), r.normal));
        if (cosTheta <= 1e-6f) return r;
        float areaPdf = 1.0f / triangleArea;
        float triangleSelectionPdf = 1.0f / (float)numTriangles;
        r.pdf = ((triangleSelectionPdf * areaPdf) * (r.distance * r.distance)) / cosTheta;
        r.emission = v3(lm.emission_color[0], lm.emission_color[1], lm.emission_color[2]) * lm.emission_power;
        r.valid = true;
    }
    return r;
}

// simple.rchit:350-385
static bool isVisibleRQ(const Ctx& cx, V3 origin, V3 target, V3 normal, uint32_t lightObject) {
    V3 toLight = normalize(target - origin);
    V3 offsetFrom = origin + (normal * 0.001f) * signf(dot(normal, toLight));
    V3 dir = target - offsetFrom;
    float dist = length(dir);
    if (dist <= 0.0f) return false;
    dir = dir / dist;
    if (cx.cnt) cx.cnt->shadow_rays++;
    bool occ = cx.cfg.use_bvh ? occluded_bvh(*cx.s, offsetFrom, dir, 0.001f, dist * 0.999f, lightObject, cx.cnt)
                              : occluded_brute(*cx.s, offsetFrom, dir, 0.001f, dist * 0.999f, lightObject);
    return !occ;
}

// simple.rchit:507-520
static inline float computeLightWeight(const Ctx& cx, V3 hitPos, uint32_t lightIdx) {
    const Scene& s = *cx.s;
    const Light& light = s.lights[lightIdx];
    const Material& lm = s.materials[s.objects[light.object_index].material_index];
    V3 lightCenter = v3(light.transform[12], light.transform[13], light.transform[14]);
    V3 toLight = lightCenter - hitPos;
    float distanceSq = dot(toLight, toLight);
    float distanceFactor = 1.0f / gmax(distanceSq, 0.001f);
    return lm.emission_power * distanceFactor;
}

// simple.rchit:472-505
static V3 evaluateLightMIS(const Ctx& cx, V3 hitPos, V3 normal, const Material& material, V3 viewDir, uint32_t lightIdx,
                           float lightSelectionPdf, uint32_t& seed) {
    V3 radiance = v3s(0.0f);
    Basis basis = createBasis(normal);
    V3 woLocal = worldToLocal(neg(viewDir), basis);
    const Light& light = cx.s->lights[lightIdx];
    LightSample ls = sampleLight(cx, lightIdx, hitPos, seed);
    if (!ls.valid) return radiance;
    V3 wiLocal = worldToLocal(ls.direction, basis);
    if (wiLocal.z > 1e-4f && isVisibleRQ(cx, hitPos, ls.position, normal, light.object_index)) {
        V3 brdf = evaluateFullBRDF(woLocal, wiLocal, material);
        float lightPdf = ls.pdf * lightSelectionPdf;
        float specularProb = getSpecularProbability(material);
        V3 hLocal = normalize(woLocal + wiLocal);
        float specularPdf = microfacetPDF(woLocal, hLocal, material.roughness);
        float diffusePdf = wiLocal.z / kPi;
        float brdfPdf = specularProb * specularPdf + (1.0f - specularProb) * diffusePdf;
        float weight = misWeightPower(lightPdf, brdfPdf);
        V3 Li = ls.emission;
        radiance = radiance + (((brdf * Li) * wiLocal.z) * weight) / gmax(lightPdf, 1e-6f);
    }
    return radiance;
}

// simple.rchit:543-583 estimateDirectLightingMIS_PowerImportance (reached through the one-line wrapper estimateDirectLightingMIS, :586-588)
static V3 estimateDirectLighting(const Ctx& cx, V3 hitPos, V3 normal, const Material& material, V3 viewDir,
                                 uint32_t currentObjectIndex, uint32_t& seed) {
    const Scene& s = *cx.s;
    int numLights = (int)s.lights.size();
    if (numLights == 0) return v3s(0.0f);
    float totalWeight = 0.0f;
    float weights[MAXLIGHTS];
    const int maxLights = numLights < MAXLIGHTS ? numLights : MAXLIGHTS;
    for (int i = 0; i < maxLights; i++) {
        weights[i] = 0.0f;  // DEFINED: the reference leaves the skipped entry uninitialised
        if (currentObjectIndex == s.lights[i].object_index) continue;
        weights[i] = computeLightWeight(cx, hitPos, (uint32_t)i);
        totalWeight += weights[i];
    }
    if (totalWeight <= 0.0f) return v3s(0.0f);
    float r1 = rnd(seed) * totalWeight;
    int selectedLight = -1;
    for (int i = 0; i < maxLights; i++) {
        r1 -= weights[i];
        if (r1 <= 0.0f) { selectedLight = i; break; }
    }
    if (selectedLight == -1) return v3s(0.0f);
    float lightPdf = weights[selectedLight] / totalWeight;
    if (cx.cnt) cx.cnt->nee++;
    return evaluateLightMIS(cx, hitPos, normal, material, viewDir, (uint32_t)selectedLight, lightPdf, seed);
}

// simple.rchit:536-541 computeLightSelectionPdf with :522-534 computeTotalLightWeight (no skip)
static float computeLightSelectionPdf(const Ctx& cx, V3 hitPos, uint32_t lightIdx) {
    float weight = computeLightWeight(cx, hitPos, lightIdx);
    int numLights = (int)cx.s->lights.size();
    const int maxLights = numLights < MAXLIGHTS ? numLights : MAXLIGHTS;
    float total = 0.0f;
    for (int i = 0; i < maxLights; i++) total += computeLightWeight(cx, hitPos, (uint32_t)i);
    if (total <= 0.0f) return 0.0f;
    return weight / total;
}

struct BSDFSample {
    V3 direction, value;
    float pdf;
    bool isSpecular;
};

// simple.rchit:403-449
static BSDFSample sampleBRDF(V3 wo, const Material& material, float rx, float ry, uint32_t& seed) {
    BSDFSample r;
    float specularWeight = getSpecularProbability(material);
    if (rnd(seed) < specularWeight) {
        V3 h = sampleGGX(rx, ry, material.roughness);
        V3 wiLocal = reflect(neg(wo), h);
        if (wiLocal.z <= 0.0f) {
            wiLocal = generateCosineWeightedDirection(rx, ry);
            r.isSpecular = false;
            r.value = diffuseValue(wo, wiLocal, material);
            r.direction = wiLocal;
        } else {
            r.isSpecular = true;
            r.value = microfacetF(wo, wiLocal, h, material);
            r.direction = wiLocal;
        }
    } else {
        r.direction = generateCosineWeightedDirection(rx, ry);
        r.isSpecular = false;
        r.value = diffuseValue(wo, r.direction, material);
    }
    V3 h = normalize(wo + r.direction);
    float specularPdf = microfacetPDF(wo, h, material.roughness);
    float diffusePdf = gmax(r.direction.z, 0.0f) / kPi;
    r.pdf = specularWeight * specularPdf + (1.0f - specularWeight) * diffusePdf;
    r.pdf = gmax(r.pdf, EPS_PDF);
    return r;
}

// simple.rchit:590-706 (main)
static void closestHitShader(const Ctx& cx, Payload& payload, const Hit& hit, V3 incomingRayDir) {
    const Scene& s = *cx.s;
    const FlatTri& ft = s.tris[hit.id];
    const uint32_t instance = ft.inst;
    const Object& object = s.objects[instance];
    const Material& mat = s.materials[object.material_index];
    const Mesh& mesh = s.meshes[object.mesh_index];
    const Vertex& v0 = mesh.v[mesh.idx[3 * ft.prim]];
    const Vertex& v1 = mesh.v[mesh.idx[3 * ft.prim + 1]];
    const Vertex& v2 = mesh.v[mesh.idx[3 * ft.prim + 2]];
    if (cx.cnt) cx.cnt->hits++;

    const float bx = (1.0f - hit.u) - hit.v, by = hit.u, bz = hit.v;
    V3 pos = (vpos(v0) * bx + vpos(v1) * by) + vpos(v2) * bz;
    V3 worldPos = xform_point(object.model, pos);
    V3 norm = normalize((vnrm(v0) * bx + vnrm(v1) * by) + vnrm(v2) * bz);
    V3 worldNrm = normalize(mul_nm(object.nrm, norm));
    bool frontFacing = dot(worldNrm, neg(incomingRayDir)) > 0.0f;
    if (!frontFacing) worldNrm = neg(worldNrm);
    bool isEmissive = mat.emission_power > 0.0f;

    bool didDirectIllumination = false;
    float p_sample_light = gclamp(mat.roughness, 0.1f, 0.9f);
    float p = rnd(payload.seed);
    if (p < p_sample_light) {
        V3 directLight = estimateDirectLighting(cx, worldPos, worldNrm, mat, incomingRayDir, instance, payload.seed);
        payload.color = payload.color + (payload.throughput * directLight) / p_sample_light;
        didDirectIllumination = true;
    }

    Basis basis = createBasis(worldNrm);
    V3 woLocal = worldToLocal(neg(incomingRayDir), basis);
    float rx = rnd(payload.seed);
    float ry = rnd(payload.seed);
    BSDFSample bs = sampleBRDF(woLocal, mat, rx, ry, payload.seed);

    if (isEmissive) {
        V3 Le = v3(mat.emission_color[0], mat.emission_color[1], mat.emission_color[2]);
        if (payload.firstBounce || payload.isSpecular) {
            payload.color = payload.color + (payload.throughput * Le) * mat.emission_power;
        } else if (!payload.didDirectIllumination) {
            uint32_t lightIdx = object.light_index;
            const Light& light = s.lights[lightIdx];
            float dist = length(worldPos - payload.previousHitPos);
            float distSq = dist * dist;
            float cosLight = gmax(0.0f, dot(worldNrm, neg(incomingRayDir)));
            V3 worldV0 = xform_point(light.transform, vpos(v0));
            V3 worldV1 = xform_point(light.transform, vpos(v1));
            V3 worldV2 = xform_point(light.transform, vpos(v2));
            float triangleArea = 0.5f * length(cross(worldV1 - worldV0, worldV2 - worldV0));
            float areaPdf = 1.0f / triangleArea;
            float triangleSelectionPdf = 1.0f / (float)light.num_triangles;
            float pdfGeo = ((triangleSelectionPdf * areaPdf) * distSq) / cosLight;
            float lightSelectionPdf = computeLightSelectionPdf(cx, worldPos, lightIdx);
            float lightPdf = lightSelectionPdf * pdfGeo;
            float brdfPdf = payload.previousBrdfPdf;
            float w = misWeightPower(brdfPdf, lightPdf);
            payload.color = payload.color + (((payload.throughput * Le) * mat.emission_power) * w) / (1.0f - payload.p_sample_light);
        }
    }
    payload.didDirectIllumination = didDirectIllumination;
    payload.p_sample_light = p_sample_light;

    if (bs.pdf > 0.0f && bs.direction.z > 0.0f) {
        payload.throughput = payload.throughput * ((bs.value * bs.direction.z) / bs.pdf);
        payload.nextDirection = localToWorld(bs.direction, basis);
        payload.hitPosition = worldPos;
        payload.hit = true;
        payload.isSpecular = bs.isSpecular;
        payload.previousBrdfPdf = bs.pdf;
        payload.previousHitPos = worldPos;
    } else {
        payload.hit = false;
    }
    payload.firstBounce = false;
}

// simple.rgen:41-53
static inline void calculateCameraRay(const float* cam, float px, float py, float W, float H, V3* origin, V3* direction) {
    const float* iv = cam + 32;  // inverse_view
    const float* ip = cam + 48;  // inverse_proj
    float ux = px / W, uy = py / H;
    float dx = ux * 2.0f - 1.0f, dy = uy * 2.0f - 1.0f;
    *origin = v3(iv[12], iv[13], iv[14]);
    V3 target = v3(((ip[0] * dx + ip[4] * dy) + ip[8]) + ip[12], ((ip[1] * dx + ip[5] * dy) + ip[9]) + ip[13],
                   ((ip[2] * dx + ip[6] * dy) + ip[10]) + ip[14]);
    V3 dl = xform_dir(iv, normalize(target));
    *direction = normalize(dl);
}

static inline bool isnan3(V3 v) { return std::isnan(v.x) || std::isnan(v.y) || std::isnan(v.z); }
static inline bool isinf3(V3 v) { return std::isinf(v.x) || std::isinf(v.y) || std::isinf(v.z); }

// simple.rgen:55-68
static bool russianRoulette(V3& throughput, uint32_t& seed, int depth) {
    if (depth < 3) return true;
    float luminance = dot(throughput, v3(0.2126f, 0.7152f, 0.0722f));
    float p = gmin(gmax(luminance, 0.05f), 0.95f);
    if (rnd(seed) > p) return false;
    throughput = throughput / p;
    return true;
}

// simple.rgen:70-137 for one pixel; returns finalColor (before accumulation)
static V3 raygenPixel(const Ctx& cx, const float* cam, const float* clear, uint32_t frame, uint32_t x, uint32_t y,
                      uint32_t W, uint32_t H) {
    uint32_t seed = tea(y * W + x, frame);
    float jx = 0.5f, jy = 0.5f;
    if (frame != 0) {  // simple.rgen:25-38 getSampleOffset
        float r1 = rnd(seed);
        float r2 = rnd(seed);
        jx = 0.5f + (r1 - 0.5f) * 0.4f;
        jy = 0.5f + (r2 - 0.5f) * 0.4f;
    }
    V3 rayOrigin, rayDirection;
    calculateCameraRay(cam, (float)x + jx, (float)y + jy, (float)W, (float)H, &rayOrigin, &rayDirection);
    Payload payload;
    payload.seed = seed;
    payload.color = v3s(0.0f);
    payload.throughput = v3s(1.0f);
    payload.firstBounce = true;
    payload.isSpecular = false;
    payload.previousBrdfPdf = 1.0f;
    payload.didDirectIllumination = false;
    payload.p_sample_light = 0.0f;
    payload.previousHitPos = v3s(0.0f);
    payload.hit = false;
    const V3 clearc = v3(clear[0], clear[1], clear[2]);
    for (int depth = 0; depth < cx.cfg.max_depth; depth++) {
        if (!russianRoulette(payload.throughput, seed, depth)) break;
        if (cx.cnt) cx.cnt->extend_rays++;
        Hit h = cx.cfg.use_bvh ? closest_bvh(*cx.s, rayOrigin, rayDirection, 0.001f, 10000.0f, cx.cnt)
                               : closest_brute(*cx.s, rayOrigin, rayDirection, 0.001f, 10000.0f);
        if (h.id == 0xFFFFFFFFu) payload.hit = false;  // simple.rmiss:26-29
        else closestHitShader(cx, payload, h, rayDirection);
        if (!payload.hit) {
            payload.color = payload.color + payload.throughput * clearc;
            break;
        }
        rayOrigin = payload.hitPosition;
        rayDirection = payload.nextDirection;
        V3 T = payload.throughput;
        if (isnan3(T) || isinf3(T) || (T.x < 0.001f && T.y < 0.001f && T.z < 0.001f)) break;
    }
    V3 fc = v3(gmin(payload.color.x, cx.cfg.clamp), gmin(payload.color.y, cx.cfg.clamp), gmin(payload.color.z, cx.cfg.clamp));
    if (isnan3(fc) || isinf3(fc)) fc = v3s(0.0f);
    return fc;
}

void render_rows(const Scene& s, const Cfg& cfg, const float* cam, const float* clear, uint32_t frame, uint32_t W,
                 uint32_t H, uint32_t y0, uint32_t y1, float* image_rgba, Counters* total) {
    Counters sum{0, 0, 0, 0, 0, 0};
#pragma omp parallel
    {
        Counters local{0, 0, 0, 0, 0, 0};
        Ctx cx{&s, cfg, total ? &local : nullptr};
#pragma omp for schedule(dynamic, 1)
        for (int y = (int)y0; y < (int)y1; y++) {
            for (uint32_t x = 0; x < W; x++) {
                V3 fc = raygenPixel(cx, cam, clear, frame, x, (uint32_t)y, W, H);
                float* px = image_rgba + 4 * ((size_t)y * W + x);
                if (frame == 0) {  // simple.rgen:127-136
                    px[0] = fc.x; px[1] = fc.y; px[2] = fc.z; px[3] = 1.0f;
                } else {
                    float a = 1.0f / (float)(frame + 1u);
                    px[0] = mixf(px[0], fc.x, a);
                    px[1] = mixf(px[1], fc.y, a);
                    px[2] = mixf(px[2], fc.z, a);
                    px[3] = 1.0f;
                }
            }
        }
#pragma omp critical
        {
            sum.extend_rays += local.extend_rays; sum.shadow_rays += local.shadow_rays; sum.nodes += local.nodes;
            sum.tris += local.tris; sum.hits += local.hits; sum.nee += local.nee;
        }
    }
    if (total) {
        total->extend_rays += sum.extend_rays; total->shadow_rays += sum.shadow_rays; total->nodes += sum.nodes;
        total->tris += sum.tris; total->hits += sum.hits; total->nee += sum.nee;
    }
}

// ---- G-buffer pass: shaders/restir.rgen:43-92, restir.rchit:36-59, restir.rmiss:18-20 ----
static inline uint16_t f32_to_f16_rne(float f) {
    uint32_t x = f2u(f);
    uint32_t sign = (x >> 16) & 0x8000u;
    uint32_t ax = x & 0x7FFFFFFFu;
    if (ax > 0x7F800000u) return (uint16_t)0x7E00u;  // NaN -> canonical quiet NaN (x86 and GPU disagree on the sign)
    if (ax == 0x7F800000u) return (uint16_t)(sign | 0x7C00u);
    if (ax >= 0x477FF000u) return (uint16_t)(sign | 0x7C00u);  // rounds to inf (>= 65520)
    if (ax < 0x33000001u) return (uint16_t)sign;                // rounds to zero (<= 2^-25)
    int exp = (int)(ax >> 23) - 127;
    uint32_t man = (ax & 0x7FFFFFu) | 0x800000u;
    int shift;
    uint32_t base;
    if (exp < -14) { shift = 13 + (-14 - exp); base = 0; }      // subnormal half
    else { shift = 13; base = (uint32_t)(exp + 15) << 10; man &= 0x7FFFFFu; }
    uint32_t q = man >> shift;
    uint32_t rem = man & ((1u << shift) - 1u);
    uint32_t half = 1u << (shift - 1);
    if (rem > half || (rem == half && (q & 1u))) q++;
    return (uint16_t)(sign | (base + q));
}
static inline uint8_t unorm8(float f) {
    float c = gclamp(f, 0.0f, 1.0f);
    if (!(c == c)) c = 0.0f;
    return (uint8_t)(int)rintf(c * 255.0f);
}

void gbuffer_rows(const Scene& s, int use_bvh, const float* cam, uint32_t W, uint32_t H, uint32_t y0, uint32_t y1,
                  uint8_t* albedo_rgba8, uint16_t* normal_rough_f16, float* position_metal_f32, uint16_t* emission_f16,
                  uint8_t* output_rgba8) {
    const float* iv = cam + 32;
    const float* ip = cam + 48;
#pragma omp parallel for schedule(dynamic, 1)
    for (int y = (int)y0; y < (int)y1; y++) {
        for (uint32_t x = 0; x < W; x++) {
            // restir.rgen:47-57: pixel corner, direction not re-normalised after the view transform
            float ux = (float)x / (float)W, uy = (float)y / (float)H;
            float dx = ux * 2.0f - 1.0f, dy = uy * 2.0f - 1.0f;
            V3 origin = v3(iv[12], iv[13], iv[14]);
            V3 target = v3(((ip[0] * dx + ip[4] * dy) + ip[8]) + ip[12], ((ip[1] * dx + ip[5] * dy) + ip[9]) + ip[13],
                           ((ip[2] * dx + ip[6] * dy) + ip[10]) + ip[14]);
            V3 dir = xform_dir(iv, normalize(target));
            Hit h = use_bvh ? closest_bvh(s, origin, dir, 0.001f, 10000.0f, nullptr)
                            : closest_brute(s, origin, dir, 0.001f, 10000.0f);
            size_t i = (size_t)y * W + x;
            if (h.id != 0xFFFFFFFFu) {
                const FlatTri& ft = s.tris[h.id];
                const Object& object = s.objects[ft.inst];
                const Material& m = s.materials[object.material_index];
                const Mesh& mesh = s.meshes[object.mesh_index];
                const Vertex& v0 = mesh.v[mesh.idx[3 * ft.prim]];
                const Vertex& v1 = mesh.v[mesh.idx[3 * ft.prim + 1]];
                const Vertex& v2 = mesh.v[mesh.idx[3 * ft.prim + 2]];
                const float bx = (1.0f - h.u) - h.v, by = h.u, bz = h.v;
                V3 pos = (vpos(v0) * bx + vpos(v1) * by) + vpos(v2) * bz;
                V3 worldPos = xform_point(object.model, pos);
                V3 norm = normalize((vnrm(v0) * bx + vnrm(v1) * by) + vnrm(v2) * bz);
                V3 worldNrm = normalize(mul_nm(object.nrm, norm));
                albedo_rgba8[4 * i + 0] = unorm8(m.albedo[0]); albedo_rgba8[4 * i + 1] = unorm8(m.albedo[1]);
                albedo_rgba8[4 * i + 2] = unorm8(m.albedo[2]); albedo_rgba8[4 * i + 3] = 255;
                normal_rough_f16[4 * i + 0] = f32_to_f16_rne(worldNrm.x); normal_rough_f16[4 * i + 1] = f32_to_f16_rne(worldNrm.y);
                normal_rough_f16[4 * i + 2] = f32_to_f16_rne(worldNrm.z); normal_rough_f16[4 * i + 3] = f32_to_f16_rne(m.roughness);
                position_metal_f32[4 * i + 0] = worldPos.x; position_metal_f32[4 * i + 1] = worldPos.y;
                position_metal_f32[4 * i + 2] = worldPos.z; position_metal_f32[4 * i + 3] = m.metallic;
                emission_f16[4 * i + 0] = f32_to_f16_rne(m.emission_color[0]); emission_f16[4 * i + 1] = f32_to_f16_rne(m.emission_color[1]);
                emission_f16[4 * i + 2] = f32_to_f16_rne(m.emission_color[2]); emission_f16[4 * i + 3] = f32_to_f16_rne(m.emission_power);
                for (int c = 0; c < 4; c++) output_rgba8[4 * i + c] = albedo_rgba8[4 * i + c];
            } else {
                const uint16_t one16 = 0x3C00;
                for (int c = 0; c < 3; c++) {
                    albedo_rgba8[4 * i + c] = 0; output_rgba8[4 * i + c] = 0;
                    normal_rough_f16[4 * i + c] = 0; emission_f16[4 * i + c] = 0; position_metal_f32[4 * i + c] = 0.0f;
                }
                albedo_rgba8[4 * i + 3] = 255; output_rgba8[4 * i + 3] = 255;
                normal_rough_f16[4 * i + 3] = one16; emission_f16[4 * i + 3] = one16; position_metal_f32[4 * i + 3] = 1.0f;
            }
        }
    }
}

// --------------------------------------------------------------------------------------------
// Rasterised G-buffer variant: shaders/restir/gbuffer.vert:22-37 and gbuffer.frag:26-39 with the fixed-function rules this
// repo states for them (no host code of the reference draws with these shaders): homogeneous edge functions over clip-space
// (x, y, w) sampled at pixel centres, no culling, perspective-correct barycentrics lambda_i = E_i / (E_0 + E_1 + E_2), samples
// with w <= 0 or depth z / w outside [0, 1] dropped, depth test LESS with ties to the lowest flat triangle id.  CPU twin of
// raytracer_b200/csrc/rt_raster.cu: the same arithmetic in the same order, written independently.
// --------------------------------------------------------------------------------------------
struct RVert { V3 wp, wn; float cx, cy, cz, cw; };

static void projView(const float* P, const float* V, float* PV) {  // (projection * view): GLSL multiplies left to right
    for (int c = 0; c < 4; c++)
        for (int r = 0; r < 4; r++)
            PV[c * 4 + r] = ((P[0 * 4 + r] * V[c * 4 + 0] + P[1 * 4 + r] * V[c * 4 + 1]) + P[2 * 4 + r] * V[c * 4 + 2]) + P[3 * 4 + r] * V[c * 4 + 3];
}

// gbuffer.vert:22-37
static RVert rasterVertex(const Object& o, const Vertex& v, const float* PV) {
    RVert r;
    V3 flippedPos = v3(v.pos[0], -v.pos[1], v.pos[2]);
    V3 flippedNormal = v3(v.normal[0], -v.normal[1], v.normal[2]);
    r.wp = xform_point(o.model, flippedPos);
    r.wn = normalize(mul_nm(o.nrm, flippedNormal));
    r.cx = ((PV[0] * r.wp.x + PV[4] * r.wp.y) + PV[8] * r.wp.z) + PV[12];
    r.cy = ((PV[1] * r.wp.x + PV[5] * r.wp.y) + PV[9] * r.wp.z) + PV[13];
    r.cz = ((PV[2] * r.wp.x + PV[6] * r.wp.y) + PV[10] * r.wp.z) + PV[14];
    r.cw = ((PV[3] * r.wp.x + PV[7] * r.wp.y) + PV[11] * r.wp.z) + PV[15];
    return r;
}

static bool rasterSample(const float a[3], const float b[3], const float c[3], const RVert v[3], uint32_t px, uint32_t py, uint32_t W,
                         uint32_t H, float lambda[3], float* depth) {
    const float X = (((float)px + 0.5f) / (float)W) * 2.0f - 1.0f;
    const float Y = (((float)py + 0.5f) / (float)H) * 2.0f - 1.0f;
    const float E0 = (a[0] * X + b[0] * Y) + c[0], E1 = (a[1] * X + b[1] * Y) + c[1], E2 = (a[2] * X + b[2] * Y) + c[2];
    if (!((E0 >= 0.0f && E1 >= 0.0f && E2 >= 0.0f) || (E0 <= 0.0f && E1 <= 0.0f && E2 <= 0.0f))) return false;
    const float S = (E0 + E1) + E2;
    if (!(S != 0.0f)) return false;
    lambda[0] = E0 / S; lambda[1] = E1 / S; lambda[2] = E2 / S;
    const float w = (lambda[0] * v[0].cw + lambda[1] * v[1].cw) + lambda[2] * v[2].cw;
    if (!(w > 0.0f)) return false;
    const float z = (lambda[0] * v[0].cz + lambda[1] * v[1].cz) + lambda[2] * v[2].cz;
    const float d = z / w;
    if (!(d >= 0.0f && d <= 1.0f)) return false;
    *depth = d;
    return true;
}

static void rasterTriangle(const Scene& s, uint32_t flat, const float* PV, RVert v[3], float a[3], float b[3], float c[3]) {
    const FlatTri& ft = s.tris[flat];
    const Object& o = s.objects[ft.inst];
    const Mesh& mesh = s.meshes[o.mesh_index];
    for (int k = 0; k < 3; k++) v[k] = rasterVertex(o, mesh.v[mesh.idx[3 * ft.prim + k]], PV);
    for (int i = 0; i < 3; i++) {
        const RVert& p = v[(i + 1) % 3];
        const RVert& q = v[(i + 2) % 3];
        a[i] = p.cy * q.cw - p.cw * q.cy;
        b[i] = p.cw * q.cx - p.cx * q.cw;
        c[i] = p.cx * q.cy - p.cy * q.cx;
    }
}

// out: five W*H*4 float images (albedo|1, normal|metallic, position|roughness, emission|isEmissive, roughness|metallic|isEmissive|0)
void raster_gbuffer(const Scene& s, const float* cam, uint32_t W, uint32_t H, float* albedo, float* normal_metal, float* position_rough,
                    float* emission, float* props) {
    float PV[16];
    projView(cam, cam + 16, PV);
    std::vector<uint64_t> vis((size_t)W * H, ~0ull);
    for (uint32_t flat = 0; flat < (uint32_t)s.tris.size(); flat++) {
        RVert v[3];
        float a[3], b[3], c[3];
        rasterTriangle(s, flat, PV, v, a, b, c);
        int x0 = 0, y0 = 0, x1 = (int)W - 1, y1 = (int)H - 1;  // conservative box: projected vertices when all lie in front
        if (v[0].cw > 1e-6f && v[1].cw > 1e-6f && v[2].cw > 1e-6f) {
            float lx = INFINITY, hx = -INFINITY, ly = INFINITY, hy = -INFINITY;
            for (int k = 0; k < 3; k++) {
                const float sx = (v[k].cx / v[k].cw * 0.5f + 0.5f) * (float)W, sy = (v[k].cy / v[k].cw * 0.5f + 0.5f) * (float)H;
                lx = fminf(lx, sx); hx = fmaxf(hx, sx); ly = fminf(ly, sy); hy = fmaxf(hy, sy);
            }
            if (lx == lx && hx == hx && ly == ly && hy == hy) {
                x0 = std::max(x0, (int)fmaxf(fminf(floorf(lx) - 2.0f, 1e9f), -1e9f)); y0 = std::max(y0, (int)fmaxf(fminf(floorf(ly) - 2.0f, 1e9f), -1e9f));
                x1 = std::min(x1, (int)fmaxf(fminf(ceilf(hx) + 2.0f, 1e9f), -1e9f)); y1 = std::min(y1, (int)fmaxf(fminf(ceilf(hy) + 2.0f, 1e9f), -1e9f));
            }
        } else if (!(v[0].cw > 0.0f) && !(v[1].cw > 0.0f) && !(v[2].cw > 0.0f)) {
            continue;  // entirely behind the eye: the interpolated w is never positive
        }
        for (int py = y0; py <= y1; py++)
            for (int px = x0; px <= x1; px++) {
                float l[3], depth;
                if (!rasterSample(a, b, c, v, (uint32_t)px, (uint32_t)py, W, H, l, &depth)) continue;
                const uint64_t key = ((uint64_t)f2u(depth) << 32) | flat;
                uint64_t& dst = vis[(size_t)py * W + px];
                if (key < dst) dst = key;
            }
    }
    for (uint32_t py = 0; py < H; py++)
        for (uint32_t px = 0; px < W; px++) {
            const size_t i = (size_t)py * W + px;
            float o[5][4] = {{0, 0, 0, 0}, {0, 0, 0, 0}, {0, 0, 0, 0}, {0, 0, 0, 0}, {0, 0, 0, 0}};
            if (vis[i] != ~0ull) {
                const uint32_t flat = (uint32_t)(vis[i] & 0xFFFFFFFFull);
                RVert v[3];
                float a[3], b[3], c[3], l[3] = {0.0f, 0.0f, 0.0f}, depth;
                rasterTriangle(s, flat, PV, v, a, b, c);
                rasterSample(a, b, c, v, px, py, W, H, l, &depth);
                V3 inPos = (v[0].wp * l[0] + v[1].wp * l[1]) + v[2].wp * l[2];
                V3 inNormal = (v[0].wn * l[0] + v[1].wn * l[1]) + v[2].wn * l[2];
                const Material& m = s.materials[s.objects[s.tris[flat].inst].material_index];
                // gbuffer.frag:26-39
                V3 n = normalize(inNormal);
                const float isEmissive = m.emission_power > 0.0f ? 1.0f : 0.0f;
                o[0][0] = m.albedo[0]; o[0][1] = m.albedo[1]; o[0][2] = m.albedo[2]; o[0][3] = 1.0f;
                o[1][0] = n.x; o[1][1] = n.y; o[1][2] = n.z; o[1][3] = m.metallic;
                o[2][0] = inPos.x; o[2][1] = inPos.y; o[2][2] = inPos.z; o[2][3] = m.roughness;
                o[3][0] = m.emission_power * m.emission_color[0]; o[3][1] = m.emission_power * m.emission_color[1];
                o[3][2] = m.emission_power * m.emission_color[2]; o[3][3] = isEmissive;
                o[4][0] = m.roughness; o[4][1] = m.metallic; o[4][2] = isEmissive; o[4][3] = 0.0f;
            }
            float* dst[5] = {albedo, normal_metal, position_rough, emission, props};
            for (int k = 0; k < 5; k++)
                if (dst[k]) memcpy(dst[k] + 4 * i, o[k], 16);
        }
}

// the two programmable stages on their own, for the comparison with gbuffer_vert.spv / gbuffer_frag.spv (tools/spirv_golden.py)
void unit_raster_vertex(const float* model16, const float* pos3, const float* normal3, const float* cam, float out[10]) {
    Object o;
    memcpy(o.model, model16, 64);
    normal_matrix(o.model, o.nrm);
    Vertex v;
    memcpy(v.pos, pos3, 12); memcpy(v.normal, normal3, 12);
    float PV[16];
    projView(cam, cam + 16, PV);
    RVert r = rasterVertex(o, v, PV);
    out[0] = r.wp.x; out[1] = r.wp.y; out[2] = r.wp.z; out[3] = r.wn.x; out[4] = r.wn.y; out[5] = r.wn.z;
    out[6] = r.cx; out[7] = r.cy; out[8] = r.cz; out[9] = r.cw;
}

// exported for unit tests of the BSDF pieces
// --------------------------------------------------------------------------------------------
// Ground truth for the ReSTIR DI path (the reference has no resampling algorithm to restate):
// direct lighting of the primary hit of restir.rgen's un-jittered corner ray, integrated by brute force —
// every triangle of every light, `spt` x `spt` stratified points each, one occlusion test per point
// (brute force or BVH, the same contract as isVisibleRQ).  Output: linear RGB, alpha 1; emissive
// surfaces add their own emission, misses return the clear colour.  No clamp.
// --------------------------------------------------------------------------------------------
void direct_reference_rows(const Scene& s, int use_bvh, const float* cam, const float* clear, uint32_t W, uint32_t H,
                           uint32_t y0, uint32_t y1, int spt, float* image_rgba) {
    const float* iv = cam + 32;
    const float* ip = cam + 48;
#pragma omp parallel for schedule(dynamic, 1)
    for (int y = (int)y0; y < (int)y1; y++) {
        for (uint32_t x = 0; x < W; x++) {
            float ux = (float)x / (float)W, uy = (float)y / (float)H;
            float dx = ux * 2.0f - 1.0f, dy = uy * 2.0f - 1.0f;
            V3 origin = v3(iv[12], iv[13], iv[14]);
            V3 target = v3(((ip[0] * dx + ip[4] * dy) + ip[8]) + ip[12], ((ip[1] * dx + ip[5] * dy) + ip[9]) + ip[13],
                           ((ip[2] * dx + ip[6] * dy) + ip[10]) + ip[14]);
            V3 dir = xform_dir(iv, normalize(target));
            Hit h = use_bvh ? closest_bvh(s, origin, dir, 0.001f, 10000.0f, nullptr) : closest_brute(s, origin, dir, 0.001f, 10000.0f);
            float* px = image_rgba + 4 * ((size_t)y * W + x);
            px[3] = 1.0f;
            if (h.id == 0xFFFFFFFFu) { px[0] = clear[0]; px[1] = clear[1]; px[2] = clear[2]; continue; }
            const FlatTri& ft = s.tris[h.id];
            const Object& object = s.objects[ft.inst];
            const Material& m = s.materials[object.material_index];
            const Mesh& mesh = s.meshes[object.mesh_index];
            const Vertex& v0 = mesh.v[mesh.idx[3 * ft.prim]];
            const Vertex& v1 = mesh.v[mesh.idx[3 * ft.prim + 1]];
            const Vertex& v2 = mesh.v[mesh.idx[3 * ft.prim + 2]];
            const float bx = (1.0f - h.u) - h.v, by = h.u, bz = h.v;
            V3 worldPos = xform_point(object.model, (vpos(v0) * bx + vpos(v1) * by) + vpos(v2) * bz);
            V3 N = normalize(mul_nm(object.nrm, normalize((vnrm(v0) * bx + vnrm(v1) * by) + vnrm(v2) * bz)));
            V3 wo_w = neg(normalize(dir));
            if (!(dot(N, wo_w) > 0.0f)) N = neg(N);
            Basis basis = createBasis(N);
            V3 wo = worldToLocal(wo_w, basis);
            double acc[3] = {0, 0, 0};
            for (size_t li = 0; li < s.lights.size(); li++) {
                const Light& L = s.lights[li];
                if (L.object_index == ft.inst) continue;
                const Object& lo = s.objects[L.object_index];
                const Material& lm = s.materials[lo.material_index];
                const Mesh& lmesh = s.meshes[lo.mesh_index];
                V3 Le = v3(lm.emission_color[0], lm.emission_color[1], lm.emission_color[2]) * lm.emission_power;
                for (uint32_t t = 0; t < L.num_triangles; t++) {
                    V3 w0 = xform_point(L.transform, vpos(lmesh.v[lmesh.idx[3 * t]]));
                    V3 w1 = xform_point(L.transform, vpos(lmesh.v[lmesh.idx[3 * t + 1]]));
                    V3 w2 = xform_point(L.transform, vpos(lmesh.v[lmesh.idx[3 * t + 2]]));
                    V3 cr = cross(w1 - w0, w2 - w0);
                    float area = 0.5f * length(cr);
                    if (!(area > 0.0f)) continue;
                    V3 ln = normalize(cr);
                    for (int a = 0; a < spt; a++)
                        for (int b = 0; b < spt; b++) {
                            float r1 = ((float)a + 0.5f) / (float)spt, r2 = ((float)b + 0.5f) / (float)spt;
                            float sq = sqrtf(r1);
                            float bu = 1.0f - sq, bv = sq * (1.0f - r2), bw = sq * r2;
                            V3 yp = (w0 * bu + w1 * bv) + w2 * bw;
                            V3 to = yp - worldPos;
                            float d2 = dot(to, to);
                            if (!(d2 > 1e-12f)) continue;
                            float d = sqrtf(d2);
                            V3 wi_w = to / d;
                            float cos_y = fabsf(dot(ln, wi_w));
                            V3 wi = worldToLocal(wi_w, basis);
                            if (!(wi.z > 1e-4f) || !(cos_y > 1e-6f)) continue;
                            // visibility exactly as isVisibleRQ sets the ray up (simple.rchit:350-371)
                            V3 from = worldPos + (N * 0.001f) * signf(dot(N, normalize(yp - worldPos)));
                            V3 sdir = yp - from;
                            float sdist = length(sdir);
                            if (!(sdist > 0.0f)) continue;
                            sdir = sdir / sdist;
                            bool occ = use_bvh ? occluded_bvh(s, from, sdir, 0.001f, sdist * 0.999f, L.object_index, nullptr)
                                               : occluded_brute(s, from, sdir, 0.001f, sdist * 0.999f, L.object_index);
                            if (occ) continue;
                            V3 f = evaluateFullBRDF(wo, wi, m);
                            float g = (wi.z * cos_y) / d2 * (area / (float)(spt * spt));
                            acc[0] += (double)(f.x * Le.x * g); acc[1] += (double)(f.y * Le.y * g); acc[2] += (double)(f.z * Le.z * g);
                        }
                }
            }
            if (m.emission_power > 0.0f) {
                acc[0] += m.emission_color[0] * m.emission_power; acc[1] += m.emission_color[1] * m.emission_power;
                acc[2] += m.emission_color[2] * m.emission_power;
            }
            px[0] = (float)acc[0]; px[1] = (float)acc[1]; px[2] = (float)acc[2];
        }
    }
}

void unit_sample_brdf(const Material& m, V3 wo, float rx, float ry, uint32_t* seed, float out[8]) {
    BSDFSample b = sampleBRDF(wo, m, rx, ry, *seed);
    out[0] = b.direction.x; out[1] = b.direction.y; out[2] = b.direction.z;
    out[3] = b.value.x; out[4] = b.value.y; out[5] = b.value.z; out[6] = b.pdf; out[7] = b.isSpecular ? 1.0f : 0.0f;
}
float unit_specular_probability(const Material& m) { return getSpecularProbability(m); }  // simple.rchit:69-75
void unit_eval_brdf(const Material& m, V3 wo, V3 wi, float out[4]) {
    V3 f = evaluateFullBRDF(wo, wi, m);
    out[0] = f.x; out[1] = f.y; out[2] = f.z;
    V3 h = normalize(wo + wi);
    float sp = getSpecularProbability(m);
    out[3] = sp * microfacetPDF(wo, h, m.roughness) + (1.0f - sp) * (gmax(wi.z, 0.0f) / kPi);
}

}  // namespace orc

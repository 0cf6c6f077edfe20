// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
// Literal restatement of the reference's host-side scene construction. Quirks are
// reproduced on purpose; each is marked.
#include "oracle_scene.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <fstream>
#include <stdexcept>

namespace orc {
using namespace om;

namespace {
const unsigned int BINS = 20;                                    // src/vk_engine.h:203
inline float iMax(float a, float b) { return a < b ? b : a; }    // src/vk_engine.h:41-43
inline float iMin(float a, float b) { return a < b ? a : b; }    // src/vk_engine.h:45-47
}  // namespace

void BoundingBox::grow(const ol::TrianglePoint& v0) {            // src/vk_engine.h:83-88
    for (int i = 0; i < 3; i++) {
        bounds[0][i] = iMin(v0.position[i], bounds[0][i]);
        bounds[1][i] = iMax(v0.position[i], bounds[1][i]);
    }
}
void BoundingBox::grow(const BoundingBox& box) {                 // src/vk_engine.h:90-93 (glm::min/max on vec4)
    for (int i = 0; i < 4; i++) {
        bounds[0][i] = gmin(bounds[0][i], box.bounds[0][i]);
        bounds[1][i] = gmax(bounds[1][i], box.bounds[1][i]);
    }
}
float BoundingBox::surfaceArea() const {                         // src/vk_engine.h:102-107 (half area)
    float x = bounds[1].x - bounds[0].x;
    float y = bounds[1].y - bounds[0].y;
    float z = bounds[1].z - bounds[0].z;
    return x * y + y * z + z * x;
}

mat4 object_matrix(const ImGuiObject& o) {                       // src/vk_engine.cpp:807-811
    return translate(o.position) * rotate(radians(o.rotation.x), vec3(1.f, 0.f, 0.f)) *
           rotate(radians(o.rotation.y), vec3(0.f, 1.f, 0.f)) * rotate(radians(o.rotation.z), vec3(0.f, 0.f, 1.f)) *
           scale(o.scale);
}

static void store_matrix(ol::RenderObject& obj, const mat4& m) {
    for (int c = 0; c < 4; ++c)
        for (int r = 0; r < 4; ++r) obj.transformMatrix[4 * c + r] = m[c][r];
}

ol::RayMaterial SceneBuilder::default_material() {               // src/vk_engine.h:69-79
    ol::RayMaterial m;
    std::memset(&m, 0, sizeof(m));
    m.albedo[0] = m.albedo[1] = m.albedo[2] = 1.f;
    m.emissionStrength = 0.f;
    m.reflectance = 0.f;
    m.ior = -1.f;
    m.albedoIndex = m.metalnessIndex = m.alphaIndex = m.bumpIndex = -1;
    return m;
}

bool SceneBuilder::read_obj(const std::string& filePath, const ImGuiObject& imGuiObj, int material) {
    // dont store the same tris, reuse bvh (:802-815)
    if (loadedObjects.count(filePath) != 0) {
        ol::RenderObject object;
        std::memset(&object, 0, sizeof(object));
        object.materialIndex = material;
        object.smoothShade = false;
        object.bvhIndex = loadedObjects.at(filePath);
        store_matrix(object, object_matrix(imGuiObj));
        object.samplerIndex = 0;
        objects.push_back(object);
        return true;
    }

    int objectTriOffset = triangles.size();
    std::ifstream fileStream;
    fileStream.open(filePath);

    bool smoothShade = false;
    bool includeUVs = false;
    std::string currentMat;
    std::string fileLine;
    std::string materialFile;
    std::vector<vec3> positions;
    std::vector<vec2> uvs;
    std::vector<vec3> normals;

    if (!fileStream.is_open()) return false;
    while (fileStream) {
        std::getline(fileStream, fileLine);
        if (fileLine.find("mtllib") != std::string::npos) {
            materialFile = fileLine.substr(7, fileLine.size() - 7);
            std::string mtlPath = filePath.substr(0, filePath.rfind("/") + 1);
            read_mtl(mtlPath + materialFile);
        }

        std::string prefix = fileLine.substr(0, fileLine.find(' '));

        if (prefix == "v") {
            int index = 2;
            vec3 position;
            for (int i = 0; i < 3; i++) {
                int nextSpace = fileLine.find(' ', index);
                position[i] = std::stof(fileLine.substr(index, nextSpace - index));
                index = nextSpace + 1;
            }
            positions.push_back(position);
        } else if (prefix == "vt") {
            vec2 uv;
            int firstSpace = fileLine.find(' ', 2);
            int secondSpace = fileLine.find(' ', firstSpace + 1);
            float u = std::stof(fileLine.substr(firstSpace + 1, secondSpace - firstSpace - 1));
            float v = std::stof(fileLine.substr(secondSpace, fileLine.length() - secondSpace));
            uv.x = u;
            uv.y = v;
            uvs.push_back(uv);
        } else if (prefix == "vn") {
            int index = 3;
            vec3 normal;
            for (int i = 0; i < 3; i++) {
                int nextSpace = fileLine.find(' ', index);
                normal[i] = std::stof(fileLine.substr(index, nextSpace - index));
                index = nextSpace + 1;
            }
            normals.push_back(normal);
        } else if (prefix == "f") {
            int index = 0;

            std::vector<uint32_t> vertexInd;
            std::vector<uint32_t> normalInd;
            std::vector<uint32_t> textureInd;

            int pointCount = 0;
            for (int i = 0; i < (int)fileLine.size() - 1; i++) {
                if (fileLine.at(i) == ' ') pointCount++;
            }

            for (int i = 0; i < pointCount; i++) {
                int space = fileLine.find(' ', index);
                int nextSpace = fileLine.find(' ', space + 1);
                std::string vertex = fileLine.substr(space + 1, nextSpace - space - (i == pointCount - 1 ? 0 : 1));

                int firstSlash = vertex.find('/');
                int secondSlash = vertex.find('/', firstSlash + 1);

                std::string vIndexStr = vertex.substr(0, firstSlash);
                if (!vIndexStr.empty()) {
                    vertexInd.push_back(std::stoi(vIndexStr) - 1);
                }

                std::string uvIndexStr = vertex.substr(firstSlash + 1, secondSlash - firstSlash - 1);
                if (!uvIndexStr.empty()) {
                    textureInd.push_back(std::stoi(uvIndexStr) - 1);
                    includeUVs = true;
                }

                if (normals.size() == 0) continue;  // quirk: `index` is not advanced (:906 vs :912)
                std::string nIndexStr = vertex.substr(secondSlash + 1, vertex.size() - secondSlash - 1);
                if (!nIndexStr.empty()) {
                    normalInd.push_back(std::stoi(nIndexStr) - 1);
                }

                index = nextSpace;
            }

            uint32_t pointIndex[4] = {0, 0, 0, 0};
            if (pointCount > 4) throw std::runtime_error("face with more than 4 vertices (UB in the reference)");

            // put uv in the vec4s (:919-935): every face owns private TrianglePoints
            for (int i = 0; i < pointCount; i++) {
                vec3 normal;
                if (normals.size() == 0) {
                    normal = vec3(0.f);
                } else {
                    normal = normals.at(normalInd.at(i));
                }
                vec2 uv = includeUVs ? uvs.at(textureInd.at(i)) : vec2{0.f, 0.f};
                vec3 p = positions.at(vertexInd.at(i));

                ol::TrianglePoint point;
                point.position[0] = p.x; point.position[1] = p.y; point.position[2] = p.z; point.position[3] = uv.x;
                point.normal[0] = normal.x; point.normal[1] = normal.y; point.normal[2] = normal.z; point.normal[3] = uv.y;

                pointIndex[i] = triPoints.size();
                triPoints.push_back(point);
            }

            ol::Triangle tri;
            std::memset(&tri, 0, sizeof(tri));   // tangent/binormal are never written by the reference (:1040-1058)
            tri.v0 = pointIndex[0];
            tri.v1 = pointIndex[1];
            tri.v2 = pointIndex[2];
            tri.frontOnly = imGuiObj.frontOnly;

            vec3 centroid = vec3(0.f);
            for (int i = 0; i < 3; i++) {
                const float* p = triPoints[pointIndex[i]].position;
                centroid.x += p[0];
                centroid.y += p[1];
                centroid.z += p[2];
            }

            triangles.push_back(tri);
            centroids.push_back(centroid / 3.f);
        } else if (prefix == "usemtl") {
            int space = fileLine.find(' ');
            std::string mat = fileLine.substr(space + 1, fileLine.size() - space - 1);
            if (currentMat.empty()) {
                currentMat = mat;
                continue;
            }
            // create object (:967-1002): one RenderObject + BVH per usemtl group
            ol::RenderObject object;
            std::memset(&object, 0, sizeof(object));
            std::string mtlPath = filePath.substr(0, filePath.rfind("/") + 1);
            object.materialIndex = loadedMaterials.at(mtlPath + materialFile + "/" + currentMat);
            store_matrix(object, object_matrix(imGuiObj));
            object.smoothShade = smoothShade;
            object.bvhIndex = bvhNodes.size();
            object.samplerIndex = imGuiObj.samplerIndex;
            objects.push_back(object);

            loadedObjects.emplace(filePath + "/" + currentMat, object.bvhIndex);

            build_bvh(triangles.size() - objectTriOffset, objectTriOffset);

            currentMat = mat;
            objectTriOffset = triangles.size();
            smoothShade = false;
        } else if (prefix == "s") {
            int smooth = fileLine.at(2) - '0';
            smoothShade = smooth == 1;
        }
    }

    ol::RenderObject object;
    std::memset(&object, 0, sizeof(object));
    std::string mtlPath = filePath.substr(0, filePath.rfind("/") + 1);
    object.materialIndex = currentMat.empty() ? material : loadedMaterials.at(mtlPath + materialFile + "/" + currentMat);
    store_matrix(object, object_matrix(imGuiObj));
    object.smoothShade = smoothShade;
    object.bvhIndex = bvhNodes.size();
    object.samplerIndex = 0;   // quirk: the end-of-file object keeps the default sampler (:1009-1019 vs :979)
    objects.push_back(object);

    loadedObjects.emplace(filePath, object.bvhIndex);

    build_bvh(triangles.size() - objectTriOffset, objectTriOffset);
    return true;
}

void SceneBuilder::add_mesh(const ol::TrianglePoint* pts, size_t nTris, const ImGuiObject& imGuiObj, int material) {
    int objectTriOffset = triangles.size();
    for (size_t t = 0; t < nTris; ++t) {
        uint32_t pointIndex[3];
        for (int i = 0; i < 3; i++) {
            pointIndex[i] = triPoints.size();
            triPoints.push_back(pts[3 * t + i]);
        }
        ol::Triangle tri;
        std::memset(&tri, 0, sizeof(tri));
        tri.v0 = pointIndex[0]; tri.v1 = pointIndex[1]; tri.v2 = pointIndex[2];
        tri.frontOnly = imGuiObj.frontOnly;
        vec3 centroid = vec3(0.f);
        for (int i = 0; i < 3; i++) {
            const float* p = triPoints[pointIndex[i]].position;
            centroid.x += p[0]; centroid.y += p[1]; centroid.z += p[2];
        }
        triangles.push_back(tri);
        centroids.push_back(centroid / 3.f);
    }
    ol::RenderObject object;
    std::memset(&object, 0, sizeof(object));
    object.materialIndex = material;
    store_matrix(object, object_matrix(imGuiObj));
    object.smoothShade = false;
    object.bvhIndex = bvhNodes.size();
    object.samplerIndex = 0;
    objects.push_back(object);
    build_bvh(triangles.size() - objectTriOffset, objectTriOffset);
}

void SceneBuilder::read_mtl(const std::string& filePath) {
    std::ifstream fileStream;
    fileStream.open(filePath);
    if (!fileStream.is_open()) return;

    std::string fileLine;
    std::string materialName;
    ol::RayMaterial currentMaterial = default_material();

    while (fileStream) {
        std::getline(fileStream, fileLine);
        if (fileLine.find("newmtl") != std::string::npos) {
            if (!materialName.empty()) {
                loadedMaterials.emplace(filePath + "/" + materialName, rayMaterials.size());
                rayMaterials.push_back(currentMaterial);
                currentMaterial = default_material();
            }
            materialName = fileLine.substr(7, fileLine.size() - 7);
            continue;
        }

        std::string mtlPath = filePath.substr(0, filePath.rfind("/") + 1);
        fileLine.erase(std::remove(fileLine.begin(), fileLine.end(), '\t'), fileLine.end());
        std::string prefix = fileLine.substr(0, fileLine.find(' '));
        auto value_of = [&]() {
            int space = fileLine.find(' ');
            return fileLine.substr(space + 1, fileLine.size() - space - 1);
        };
        if (prefix == "Ka" || prefix == "Kd") {   // quirk Q18: both multiply into albedo
            int space1 = fileLine.find(' ');
            int space2 = fileLine.find(' ', space1 + 1);
            int space3 = fileLine.find(' ', space2 + 1);
            std::string r = fileLine.substr(space1 + 1, space2 - space1 - 1);
            std::string g = fileLine.substr(space2 + 1, space3 - space2 - 1);
            std::string b = fileLine.substr(space3 + 1, fileLine.size() - space3 - 1);
            currentMaterial.albedo[0] *= std::stof(r);
            currentMaterial.albedo[1] *= std::stof(g);
            currentMaterial.albedo[2] *= std::stof(b);
        } else if (prefix == "map_Ka" || prefix == "map_Kd") {
            textureFiles.push_back(mtlPath + value_of());
            currentMaterial.albedoIndex = texturesUsed;
            texturesUsed++;
        } else if (prefix == "map_Ks") {
            textureFiles.push_back(mtlPath + value_of());
            currentMaterial.metalnessIndex = texturesUsed;
            texturesUsed++;
        } else if (prefix == "map_d") {
            textureFiles.push_back(mtlPath + value_of());
            currentMaterial.alphaIndex = texturesUsed;
            texturesUsed++;
        } else if (prefix == "map_bump") {        // quirk Q18: lower-case only
            textureFiles.push_back(mtlPath + value_of());
            currentMaterial.bumpIndex = texturesUsed;
            texturesUsed++;
        }
        // Ni / d are parsed and ignored by the reference (:1101-1108)
    }

    loadedMaterials.emplace(filePath + "/" + materialName, rayMaterials.size());
    rayMaterials.push_back(currentMaterial);
}

void SceneBuilder::build_bvh(int size, int triIndex) {           // src/vk_engine.cpp:1169-1194
    if (size <= 0) throw std::runtime_error("build_bvh on an empty triangle range (UB in the reference)");
    nodesUsed++;
    int offset = bvhNodes.size();
    ol::BVHNode zero;
    std::memset(&zero, 0, sizeof(zero));
    bvhNodes.resize(bvhNodes.size() + (size * 2 - 1), zero);
    ol::BVHNode& root = bvhNodes[offset];
    root.index = triIndex;
    root.triCount = size;

    update_bvh_bounds(offset);
    subdivide_bvh(offset, 0);

    bvhNodes.resize(nodesUsed);
}

void SceneBuilder::update_bvh_bounds(uint32_t index) {           // src/vk_engine.cpp:1196-1210
    ol::BVHNode& node = bvhNodes[index];
    BoundingBox box;
    for (uint32_t i = 0; i < node.triCount; i++) {
        const ol::Triangle& leafTri = triangles[node.index + i];
        box.grow(triPoints[leafTri.v0]);
        box.grow(triPoints[leafTri.v1]);
        box.grow(triPoints[leafTri.v2]);
    }
    node.boundsX[0] = box.bounds[0].x; node.boundsX[1] = box.bounds[1].x;
    node.boundsY[0] = box.bounds[0].y; node.boundsY[1] = box.bounds[1].y;
    node.boundsZ[0] = box.bounds[0].z; node.boundsZ[1] = box.bounds[1].z;
}

void SceneBuilder::subdivide_bvh(uint32_t index, uint32_t depth) {  // src/vk_engine.cpp:1212-1277
    ol::BVHNode& node = bvhNodes[index];

    if (node.triCount <= 2 || depth >= 64) return;

    int axis = 0;
    float splitPos = 0.f;
    float bestCost = find_bvh_split_plane(node, axis, splitPos);

    BoundingBox parent;
    parent.bounds[0] = vec4(node.boundsX[0], node.boundsY[0], node.boundsZ[0], 0.f);
    parent.bounds[1] = vec4(node.boundsX[1], node.boundsY[1], node.boundsZ[1], 0.f);
    float noSplitCost = node.triCount * parent.surfaceArea();
    if (bestCost >= noSplitCost) return;

    // partition the triangles
    int i = node.index;
    int j = i + node.triCount - 1;
    while (i <= j) {
        vec3 centroid = centroids[i];
        if (centroid[axis] < splitPos) {
            i++;
        } else {
            std::swap(triangles[i], triangles[j]);
            std::swap(centroids[i], centroids[j]);
            j--;
        }
    }

    // if one side has all tris, abort
    int triIndex = node.index;
    int leftCount = i - triIndex;
    if (leftCount == 0 || leftCount == (int)node.triCount) return;

    node.index = nodesUsed;
    nodesUsed += 2;
    bvhNodes[node.index].index = triIndex;
    bvhNodes[node.index].triCount = leftCount;
    bvhNodes[node.index + 1].index = i;
    bvhNodes[node.index + 1].triCount = node.triCount - leftCount;

    node.triCount = 0;
    update_bvh_bounds(node.index);
    update_bvh_bounds(node.index + 1);

    subdivide_bvh(node.index, depth + 1);
    subdivide_bvh(node.index + 1, depth + 1);
}

float SceneBuilder::find_bvh_split_plane(const ol::BVHNode& node, int& axis, float& splitPos) {  // :1279-1337
    struct BVHBin { BoundingBox box; uint32_t triCount = 0; };   // src/vk_engine.h:191-194
    float bestCost = 1e30f;
    for (int a = 0; a < 3; a++) {
        float min = 1e30f;
        float max = -1e30f;
        for (uint32_t i = 0; i < node.triCount; i++) {
            min = iMin(min, centroids[node.index + i][a]);
            max = iMax(max, centroids[node.index + i][a]);
        }

        if (min == max) continue;

        // populate bins
        BVHBin bins[BINS];
        float scale = BINS / (max - min);
        for (uint32_t i = 0; i < node.triCount; i++) {
            const ol::Triangle& tri = triangles[node.index + i];
            int binIndex = iMin(BINS - 1, std::floor((centroids[node.index + i][a] - min) * scale));
            bins[binIndex].triCount++;
            bins[binIndex].box.grow(triPoints[tri.v0]);
            bins[binIndex].box.grow(triPoints[tri.v1]);
            bins[binIndex].box.grow(triPoints[tri.v2]);
        }

        float leftArea[BINS - 1];
        float rightArea[BINS - 1];
        float leftCount[BINS - 1];
        float rightCount[BINS - 1];
        BoundingBox leftBox;
        BoundingBox rightBox;
        int leftSum = 0;
        int rightSum = 0;

        for (unsigned i = 0; i < BINS - 1; i++) {
            leftSum += bins[i].triCount;
            leftCount[i] = leftSum;
            leftBox.grow(bins[i].box);
            leftArea[i] = leftBox.surfaceArea();             // scene_interior_cost == surfaceArea (:1340-1341)
            rightSum += bins[BINS - 1 - i].triCount;
            rightCount[BINS - 2 - i] = rightSum;
            rightBox.grow(bins[BINS - 1 - i].box);
            rightArea[i] = rightBox.surfaceArea();           // quirk Q14: both slots written every iteration
            rightArea[BINS - 2 - i] = rightBox.surfaceArea();
        }

        scale = (max - min) / BINS;
        for (unsigned i = 0; i < BINS - 1; i++) {
            float cost = leftCount[i] * leftArea[i] + rightCount[i] * rightArea[i];
            if (cost < bestCost) {
                axis = a;
                splitPos = min + scale * (i + 1);
                bestCost = cost;
            }
        }
    }
    return bestCost;
}

void SceneBuilder::push_default_materials() {                    // src/vk_engine.cpp:689-722
    ol::RayMaterial dielectric = default_material();
    dielectric.ior = 2.f;
    ol::RayMaterial mirror = default_material();
    mirror.reflectance = 1.f;
    ol::RayMaterial white = default_material();
    ol::RayMaterial red = default_material();
    red.albedo[0] = 1.f; red.albedo[1] = 0.f; red.albedo[2] = 0.f;
    ol::RayMaterial green = default_material();
    green.albedo[0] = 0.f; green.albedo[1] = 1.f; green.albedo[2] = 0.f;
    ol::RayMaterial li = default_material();
    li.emissionColor[0] = li.emissionColor[1] = li.emissionColor[2] = 1.f;
    li.albedo[0] = li.albedo[1] = li.albedo[2] = 0.f;
    li.emissionStrength = 2.4f;

    rayMaterials.push_back(white);
    rayMaterials.push_back(red);
    rayMaterials.push_back(green);
    rayMaterials.push_back(li);
    rayMaterials.push_back(mirror);
    rayMaterials.push_back(dielectric);
}

void SceneBuilder::cornell_box(const std::string& dir) {         // src/vk_engine.cpp:638-678
    ImGuiObject light;
    light.frontOnly = true;
    light.name = "light";
    light.position = vec3(0.f, -1.5f, 0.f);
    light.scale = vec3(1.f);
    read_obj(dir + "light2.obj", light, 3);

    ImGuiObject plane;
    plane.frontOnly = true;
    plane.name = "bottom";
    plane.position = vec3(0.f, 0.5f, 0.f);
    plane.rotation = vec3(0.f);
    read_obj(dir + "plane.obj", plane, 0);

    plane.name = "left";
    plane.position = vec3(-1.f, -0.5f, 0.f);
    plane.rotation = vec3(90.f, 0.f, 90.f);
    read_obj(dir + "plane.obj", plane, 2);

    plane.name = "right";
    plane.position = vec3(1.f, -0.5f, 0.f);
    plane.rotation = vec3(90.f, 0.f, -90.f);
    read_obj(dir + "plane.obj", plane, 1);

    plane.name = "top";
    plane.position = vec3(0.f, -1.5f, 0.f);
    plane.rotation = vec3(0.f, 0.f, 0.f);
    read_obj(dir + "ceiling.obj", plane, 0);

    plane.name = "back";
    plane.position = vec3(0.f, -0.5f, 1.f);
    plane.rotation = vec3(90.f, 0.f, 0.f);
    plane.scale = vec3(1.f);
    read_obj(dir + "plane.obj", plane, 0);

    plane.name = "front";
    plane.position = vec3(0.f, -0.5f, -1.f);
    plane.rotation = vec3(-90.f, 0.f, 0.f);
    read_obj(dir + "plane.obj", plane, 0);
}

void SceneBuilder::prepare_default_scene(const std::string& dir) {  // src/vk_engine.cpp:680-758
    ol::Sphere zero;
    std::memset(&zero, 0, sizeof(zero));
    spheres.assign(10, zero);                                    // MAX_SPHERES zero spheres (:682)
    push_default_materials();

    ImGuiObject model;
    model.name = "cube";
    model.scale = vec3(0.25f);
    model.samplerIndex = 1;
    model.rotation = vec3(0.f, -30.f, 0.f);
    model.position = vec3(-0.4f, 0.25f, -0.45f);
    read_obj(dir + "cube.obj", model, 0);

    model.name = "cube2";
    model.scale = vec3(0.3f, 0.7f, 0.3f);
    model.samplerIndex = 1;
    model.rotation = vec3(0.f, 30.f, 0.f);
    model.position = vec3(0.4f, -0.2f, 0.45f);
    read_obj(dir + "cube.obj", model, 0);

    cornell_box(dir);
}

void set_camera_angles(ol::PushConstants& pc, float ax, float ay, float az) {  // src/vk_engine.cpp:1633-1653
    float thetaX = radians(ax);
    float thetaY = radians(ay);
    float thetaZ = radians(az);
    mat3 rotX, rotY, rotZ;
    rotX[0] = vec3(1, 0, 0); rotX[1] = vec3(0, std::cos(thetaX), -std::sin(thetaX)); rotX[2] = vec3(0, std::sin(thetaX), std::cos(thetaX));
    rotY[0] = vec3(std::cos(thetaY), 0, std::sin(thetaY)); rotY[1] = vec3(0, 1, 0); rotY[2] = vec3(-std::sin(thetaY), 0, std::cos(thetaY));
    rotZ[0] = vec3(std::cos(thetaZ), -std::sin(thetaZ), 0); rotZ[1] = vec3(std::sin(thetaZ), std::cos(thetaZ), 0); rotZ[2] = vec3(0, 0, 1);
    mat4 rot(rotY * rotX * rotZ);
    for (int c = 0; c < 4; ++c)
        for (int r = 0; r < 4; ++r) pc.camInfo.cameraRotation[4 * c + r] = rot[c][r];
}

ol::PushConstants default_push_constants(int W, int H, uint32_t sphereCount, uint32_t objectCount) {
    ol::PushConstants pc;
    std::memset(&pc, 0, sizeof(pc));
    // CameraInfo defaults, src/vk_engine.h:145-151
    pc.camInfo.pos[0] = 0.f; pc.camInfo.pos[1] = -0.5f; pc.camInfo.pos[2] = -3.5f;
    pc.camInfo.nearPlane = 0.1f;
    pc.camInfo.aspectRatio = W / (float)H;                        // :1631
    pc.camInfo.fov = 50.f;
    set_camera_angles(pc, 4.f, 0.f, 0.f);                         // cameraAngles, src/vk_engine.h:325
    // EnvironmentData defaults, src/vk_engine.h:153-158
    float h[4] = {0.986f, 1.f, 0.902f, 1000.f}, z[4] = {0.265f, 0.595f, 0.887f, 10.f};
    std::memcpy(pc.environment.horizonColor, h, 16);
    std::memcpy(pc.environment.zenithColor, z, 16);
    pc.environment.groundColor[0] = pc.environment.groundColor[1] = pc.environment.groundColor[2] = 0.431f;
    vec3 ld = normalize(vec3(2.f, 0.8f, -3.f));
    pc.environment.lightDir[0] = ld.x; pc.environment.lightDir[1] = ld.y; pc.environment.lightDir[2] = ld.z;
    pc.environment.lightDir[3] = 0.f;
    // RayTracerData defaults, src/vk_engine.h:160-171
    pc.rayTraceParams.progressive = 0;
    pc.rayTraceParams.singleRender = 0;
    pc.rayTraceParams.debug = -1;
    pc.rayTraceParams.raysPerPixel = 1;
    pc.rayTraceParams.bounceLimit = 8;
    pc.rayTraceParams.sphereCount = sphereCount;                  // :1655
    pc.rayTraceParams.objectCount = objectCount;                  // :1656
    pc.rayTraceParams.triangleCap = 50;
    pc.rayTraceParams.boxCap = 200;
    pc.rayTraceParams.sampleLimit = 10;
    pc.frameCount = 0;
    return pc;
}

}  // namespace orc

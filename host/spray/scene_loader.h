// scene_loader.h — `.spray` scene description parser (reference src/io/scene_loader.h:33-117,
// scene_loader.cc:40-358) and the Domain / Light / Bsdf descriptors it fills
// (src/render/domain.h:31-46, src/render/light.h:41-89, src/render/reflection.h:241-364).
//
// Tokens: `domain`, `file <ply>`, `mtl diffuse|mirror r g b | glass|transmission etaA etaB`,
// `bound minx miny minz maxx maxy maxz`, `scale x y z`, `rotate x|y|z <degrees>`, `translate x y z`,
// `face <n>`, `vertex <n>`, `light point px py pz r g b | diffuse r g b`, `#` comments.
// Transform tokens compose in file order with glm semantics; world_aabb = transform * {min, max} corners
// only (scene_loader.cc:352-357 — wrong under rotation, reproduced on purpose).
#pragma once
#include <cstddef>
#include <string>
#include <vector>

#include "spray/vecmath.h"
#include "spray_gpu.h"

namespace spray {

// The reference's Bsdf class hierarchy carries only these parameters; shading itself runs on the GPU.
struct Bsdf {
  int type = SPRAYGPU_MTL_DIFFUSE;  // SPRAYGPU_MTL_*
  float params[3] = {1.f, 1.f, 1.f};  // albedo | reflectance | etaA, etaB
  bool isDelta() const { return type != SPRAYGPU_MTL_DIFFUSE; }
};

struct Light {
  int type = SPRAYGPU_LIGHT_POINT;  // SPRAYGPU_LIGHT_*
  vec3 position;                    // point lights
  vec3 radiance;
  bool isAreaLight() const { return type == SPRAYGPU_LIGHT_DIFFUSE; }
};

struct Domain {
  unsigned id = 0;
  std::size_t num_vertices = 0;
  std::size_t num_faces = 0;
  std::string filename;
  Aabb object_aabb;
  Aabb world_aabb;
  mat4 transform;
  Bsdf bsdf;             // the reference owns a heap Bsdf* (domain.h:34); a value suffices here
  bool has_bsdf = false;
};

class SceneLoader {
 public:
  // lights_out may be nullptr only if the file has no `light` lines (scene_loader.h:96-101)
  void load(const std::string& filename, const std::string& ply_path, std::vector<Domain>* domains_out,
            std::vector<Light*>* lights_out);

 private:
  void parseLineTokens(const std::string& ply_path, const std::vector<std::string>& tokens);
  Domain& currentDomain();

  int domain_id_ = -1;
  int light_id_ = 0;
  std::vector<Domain>* domains_ = nullptr;
  std::vector<Light*>* lights_ = nullptr;
};

}  // namespace spray

// scene_loader.cc — see spray/scene_loader.h.  Semantics follow reference src/io/scene_loader.cc.
#include "spray/scene_loader.h"

#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <sstream>

#include "spray/check.h"

namespace spray {

namespace {

// strtok(line, " ") semantics: split on runs of spaces only (tabs are NOT separators in the reference)
std::vector<std::string> splitSpaces(const std::string& line) {
  std::vector<std::string> out;
  std::size_t i = 0;
  while (i < line.size()) {
    while (i < line.size() && line[i] == ' ') ++i;
    if (i >= line.size()) break;
    std::size_t j = i;
    while (j < line.size() && line[j] != ' ') ++j;
    out.push_back(line.substr(i, j - i));
    i = j;
  }
  return out;
}

float tof(const std::string& s) { return (float)std::atof(s.c_str()); }  // atof (double) -> glm::vec3 float

}  // namespace

Domain& SceneLoader::currentDomain() {
  SPRAY_CHECK(domain_id_ > -1, "token before the first `domain`");
  SPRAY_CHECK((std::size_t)domain_id_ < domains_->size(), "");
  return (*domains_)[(std::size_t)domain_id_];
}

void SceneLoader::parseLineTokens(const std::string& ply_path, const std::vector<std::string>& t) {
  const std::string& tag = t[0];
  if (tag[0] == '#') return;
  if (tag == "domain") {
    ++domain_id_;
    Domain& d = currentDomain();
    d.id = (unsigned)domain_id_;
    d.transform.setIdentity();
  } else if (tag == "file") {
    SPRAY_CHECK(t.size() == 2, "file <name>");
    currentDomain().filename = ply_path.empty() ? t[1] : ply_path + "/" + t[1];
  } else if (tag == "mtl") {
    Domain& d = currentDomain();
    SPRAY_CHECK(t.size() >= 2, "mtl <type> ...");
    if (t[1] == "diffuse" || t[1] == "mirror") {
      SPRAY_CHECK(t.size() == 5, "mtl " << t[1] << " r g b");
      d.bsdf.type = t[1] == "diffuse" ? SPRAYGPU_MTL_DIFFUSE : SPRAYGPU_MTL_MIRROR;
      for (int k = 0; k < 3; ++k) d.bsdf.params[k] = tof(t[2 + k]);
    } else if (t[1] == "glass" || t[1] == "transmission") {
      SPRAY_CHECK(t.size() == 4, "mtl " << t[1] << " etaA etaB");
      d.bsdf.type = t[1] == "glass" ? SPRAYGPU_MTL_GLASS : SPRAYGPU_MTL_TRANSMISSION;
      d.bsdf.params[0] = tof(t[2]); d.bsdf.params[1] = tof(t[3]); d.bsdf.params[2] = 0.f;
    } else {
      SPRAY_FATAL("unknown material type " << t[1]);
    }
    d.has_bsdf = true;
  } else if (tag == "bound") {
    SPRAY_CHECK(t.size() == 7, "bound minx miny minz maxx maxy maxz");
    Domain& d = currentDomain();
    d.object_aabb.bounds[0] = vec3(tof(t[1]), tof(t[2]), tof(t[3]));
    d.object_aabb.bounds[1] = vec3(tof(t[4]), tof(t[5]), tof(t[6]));
  } else if (tag == "scale") {
    SPRAY_CHECK(t.size() == 4, "scale x y z");
    Domain& d = currentDomain();
    d.transform = scale(d.transform, vec3(tof(t[1]), tof(t[2]), tof(t[3])));
  } else if (tag == "rotate") {
    SPRAY_CHECK(t.size() == 3, "rotate <axis> <degrees>");
    vec3 axis;
    if (t[1] == "x") axis = vec3(1.f, 0.f, 0.f);
    else if (t[1] == "y") axis = vec3(0.f, 1.f, 0.f);
    else if (t[1] == "z") axis = vec3(0.f, 0.f, 1.f);
    else SPRAY_FATAL("invalid axis name " << t[1]);
    Domain& d = currentDomain();
    d.transform = rotate(d.transform, (float)radians(std::atof(t[2].c_str())), axis);
  } else if (tag == "translate") {
    SPRAY_CHECK(t.size() == 4, "translate x y z");
    Domain& d = currentDomain();
    d.transform = translate(d.transform, vec3(tof(t[1]), tof(t[2]), tof(t[3])));
  } else if (tag == "face") {
    SPRAY_CHECK(t.size() == 2, "face <n>");
    currentDomain().num_faces = std::stoul(t[1]);
  } else if (tag == "vertex") {
    SPRAY_CHECK(t.size() == 2, "vertex <n>");
    currentDomain().num_vertices = std::stoul(t[1]);
  } else if (tag == "light") {
    SPRAY_CHECK(t.size() >= 2, "light <type> ...");
    Light* l = new Light;
    if (t[1] == "point") {
      SPRAY_CHECK(t.size() == 8, "light point px py pz r g b");
      l->type = SPRAYGPU_LIGHT_POINT;
      l->position = vec3(tof(t[2]), tof(t[3]), tof(t[4]));
      l->radiance = vec3(tof(t[5]), tof(t[6]), tof(t[7]));
    } else if (t[1] == "diffuse") {
      SPRAY_CHECK(t.size() == 5, "light diffuse r g b");
      l->type = SPRAYGPU_LIGHT_DIFFUSE;
      l->radiance = vec3(tof(t[2]), tof(t[3]), tof(t[4]));
    } else {
      SPRAY_FATAL("unknown light source " << t[1]);
    }
    SPRAY_CHECK(lights_ != nullptr, "");
    SPRAY_CHECK((std::size_t)light_id_ < lights_->size(), "");
    (*lights_)[(std::size_t)light_id_++] = l;
  } else {
    SPRAY_FATAL("unknown tag name " << tag);
  }
}

void SceneLoader::load(const std::string& filename, const std::string& ply_path, std::vector<Domain>* domains_out,
                       std::vector<Light*>* lights_out) {
  domain_id_ = -1;
  light_id_ = 0;
  SPRAY_CHECK(domains_out != nullptr, "");
  SPRAY_CHECK(domains_out->size() == 0, "");
  if (lights_out) SPRAY_CHECK(lights_out->size() == 0, "");
  domains_ = domains_out;
  lights_ = lights_out;

  std::ifstream infile(filename);
  SPRAY_CHECK(infile.is_open(), "unable to open input file " << filename);
  std::vector<std::vector<std::string>> lines;
  int ndomains = 0, nlights = 0;
  for (std::string line; std::getline(infile, line);) {
    std::vector<std::string> tokens = splitSpaces(line);
    if (tokens.empty()) continue;
    if (tokens[0] == "domain") ++ndomains;
    else if (tokens[0] == "light") ++nlights;
    lines.push_back(std::move(tokens));
  }
  // countAndAllocate (scene_loader.cc:281-310)
  SPRAY_CHECK(ndomains > 0, "no domains in " << filename);
  domains_->resize((std::size_t)ndomains);
  if (nlights) {
    SPRAY_CHECK(lights_ != nullptr, "");
    lights_->resize((std::size_t)nlights, nullptr);
  } else {
    std::cout << "[WARNING] no lights detected\n";
  }
  std::cout << "[INFO] number of domains: " << ndomains << "\n";
  std::cout << "[INFO] number of lights: " << nlights << "\n";

  for (const auto& tokens : lines) parseLineTokens(ply_path, tokens);

  // world bounds: transform * min / max corners only (scene_loader.cc:352-357)
  for (Domain& d : *domains_out) {
    const vec4 mn = mul(d.transform, d.object_aabb.bounds[0].x, d.object_aabb.bounds[0].y, d.object_aabb.bounds[0].z, 1.0f);
    const vec4 mx = mul(d.transform, d.object_aabb.bounds[1].x, d.object_aabb.bounds[1].y, d.object_aabb.bounds[1].z, 1.0f);
    d.world_aabb.bounds[0] = vec3(mn.x, mn.y, mn.z);
    d.world_aabb.bounds[1] = vec3(mx.x, mx.y, mx.z);
  }
}

}  // namespace spray

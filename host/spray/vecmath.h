// vecmath.h — the few glm operations the host side needs, restated with glm 0.9.9.3's operation order
// (reference deps/glm) so that scene transforms, world bounds and the camera frame come out bit-identical
// to what SpRay computes.  The host library is built with -ffp-contract=off.
#pragma once
#include <cmath>
#include <cstring>

namespace spray {

struct vec3 {
  float x, y, z;
  vec3() : x(0.f), y(0.f), z(0.f) {}
  vec3(float a, float b, float c) : x(a), y(b), z(c) {}
  float& operator[](int i) { return (&x)[i]; }
  const float& operator[](int i) const { return (&x)[i]; }
};
inline vec3 operator+(vec3 a, vec3 b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(vec3 a, vec3 b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator*(vec3 a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
inline vec3 operator*(float s, vec3 a) { return vec3(s * a.x, s * a.y, s * a.z); }
// glm::dot (detail/func_geometric.inl:46-54): tmp = a*b; tmp.x + tmp.y + tmp.z
inline float dot(vec3 a, vec3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
// glm::cross (func_geometric.inl:60-70)
inline vec3 cross(vec3 x, vec3 y) {
  return vec3(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y);
}
inline float length(vec3 v) { return std::sqrt(dot(v, v)); }
// glm::normalize = v * inversesqrt(dot(v,v)), inversesqrt = 1/sqrt (func_geometric.inl:82-89,
// func_exponential.inl:134-139)
inline vec3 normalize(vec3 v) { return v * (1.0f / std::sqrt(dot(v, v))); }

struct vec4 { float x, y, z, w; };

// column-major 4x4, m[c][r] as glm
struct mat4 {
  float m[4][4];
  mat4() { setIdentity(); }
  void setIdentity() {
    std::memset(m, 0, sizeof(m));
    m[0][0] = m[1][1] = m[2][2] = m[3][3] = 1.0f;
  }
  bool isIdentity() const {
    mat4 i;
    return std::memcmp(m, i.m, sizeof(m)) == 0;  // glm operator!= compares element-wise; +0/-0 aside
  }
  const float* data() const { return &m[0][0]; }
};

// mat4 * vec4 (detail/type_mat4x4.inl operator*, the non-SIMD path):
//   (m[0]*v.x + m[1]*v.y) + (m[2]*v.z + m[3]*v.w)
inline vec4 mul(const mat4& M, float x, float y, float z, float w) {
  vec4 r;
  float* o = &r.x;
  for (int k = 0; k < 4; ++k) o[k] = (M.m[0][k] * x + M.m[1][k] * y) + (M.m[2][k] * z + M.m[3][k] * w);
  return r;
}

// glm::translate (ext/matrix_transform.inl:10-15): Result[3] = m[0]*v[0] + m[1]*v[1] + m[2]*v[2] + m[3]
inline mat4 translate(const mat4& M, vec3 v) {
  mat4 R = M;
  for (int k = 0; k < 4; ++k) R.m[3][k] = ((M.m[0][k] * v.x + M.m[1][k] * v.y) + M.m[2][k] * v.z) + M.m[3][k];
  return R;
}
// glm::scale (ext/matrix_transform.inl): Result[i] = m[i] * v[i], Result[3] = m[3]
inline mat4 scale(const mat4& M, vec3 v) {
  mat4 R = M;
  for (int k = 0; k < 4; ++k) { R.m[0][k] = M.m[0][k] * v.x; R.m[1][k] = M.m[1][k] * v.y; R.m[2][k] = M.m[2][k] * v.z; }
  return R;
}
// glm::rotate (ext/matrix_transform.inl:17-48)
inline mat4 rotate(const mat4& M, float angle, vec3 v) {
  const float a = angle, c = std::cos(a), s = std::sin(a);
  const vec3 axis = normalize(v);
  const vec3 temp = (1.0f - c) * axis;
  float R[3][3];
  R[0][0] = c + temp[0] * axis[0];
  R[0][1] = temp[0] * axis[1] + s * axis[2];
  R[0][2] = temp[0] * axis[2] - s * axis[1];
  R[1][0] = temp[1] * axis[0] - s * axis[2];
  R[1][1] = c + temp[1] * axis[1];
  R[1][2] = temp[1] * axis[2] + s * axis[0];
  R[2][0] = temp[2] * axis[0] + s * axis[1];
  R[2][1] = temp[2] * axis[1] - s * axis[0];
  R[2][2] = c + temp[2] * axis[2];
  mat4 out;
  for (int k = 0; k < 4; ++k) {
    out.m[0][k] = (M.m[0][k] * R[0][0] + M.m[1][k] * R[0][1]) + M.m[2][k] * R[0][2];
    out.m[1][k] = (M.m[0][k] * R[1][0] + M.m[1][k] * R[1][1]) + M.m[2][k] * R[1][2];
    out.m[2][k] = (M.m[0][k] * R[2][0] + M.m[1][k] * R[2][1]) + M.m[2][k] * R[2][2];
    out.m[3][k] = M.m[3][k];
  }
  return out;
}
// glm::radians<double>: degrees * 0.01745329251994329576923690768489
inline double radians(double deg) { return deg * 0.01745329251994329576923690768489; }

// axis-aligned box — reference src/render/aabb.h (bounds[0] = min, bounds[1] = max)
struct Aabb {
  vec3 bounds[2];
  Aabb() { reset(); }
  void reset() {
    bounds[0] = vec3(INFINITY, INFINITY, INFINITY);
    bounds[1] = vec3(-INFINITY, -INFINITY, -INFINITY);
  }
  bool isValid() const {
    return bounds[0].x <= bounds[1].x && bounds[0].y <= bounds[1].y && bounds[0].z <= bounds[1].z;
  }
  void merge(const Aabb& b) {
    for (int a = 0; a < 3; ++a) {
      bounds[0][a] = std::fmin(bounds[0][a], b.bounds[0][a]);
      bounds[1][a] = std::fmax(bounds[1][a], b.bounds[1][a]);
    }
  }
  vec3 getExtent() const { return bounds[1] - bounds[0]; }
  // aabb.cc:133-137: 0.5 (double) * (float sum), narrowed to float
  vec3 getCenter() const {
    return vec3((float)(0.5 * (bounds[0].x + bounds[1].x)), (float)(0.5 * (bounds[0].y + bounds[1].y)),
                (float)(0.5 * (bounds[0].z + bounds[1].z)));
  }
};

}  // namespace spray

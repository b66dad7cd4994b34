// ply_loader.cc — see spray/ply_loader.h.  Behaviour follows reference src/io/ply_loader.cc; the
// implementation is a bulk read + in-memory decode.
#include "spray/ply_loader.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <sstream>

#include "spray/check.h"

namespace spray {

namespace {

enum { P_NONE = 0, P_X, P_Y, P_Z, P_RED, P_GREEN, P_BLUE };

// whole file -> memory in one read
std::vector<char> slurp(const std::string& filename) {
  FILE* f = std::fopen(filename.c_str(), "rb");
  SPRAY_CHECK(f != nullptr, filename);
  std::fseek(f, 0, SEEK_END);
  const long sz = std::ftell(f);
  std::fseek(f, 0, SEEK_SET);
  std::vector<char> buf((std::size_t)(sz > 0 ? sz : 0));
  if (sz > 0) {
    const std::size_t got = std::fread(buf.data(), 1, (std::size_t)sz, f);
    SPRAY_CHECK(got == (std::size_t)sz, "short read on " << filename);
  }
  std::fclose(f);
  return buf;
}

// next '\n'-terminated line starting at off (std::getline semantics: the '\n' is consumed, a trailing
// '\r' is NOT stripped — the reference does not strip it either)
bool nextLine(const char* buf, std::size_t size, std::size_t* off, std::string* line) {
  if (*off >= size) return false;
  const char* b = buf + *off;
  const char* e = (const char*)std::memchr(b, '\n', size - *off);
  if (!e) { line->assign(b, size - *off); *off = size; return true; }
  line->assign(b, (std::size_t)(e - b));
  *off = (std::size_t)(e - buf) + 1;
  return true;
}

}  // namespace

int PlyLoader::dataType(const std::string& s) {
  if (s == "char") return kCHAR;
  if (s == "uchar") return kUCHAR;
  if (s == "short") return kSHORT;
  if (s == "ushort") return kUSHORT;
  if (s == "int") return kINT;
  if (s == "uint") return kUINT;
  if (s == "float") return kFLOAT;
  if (s == "double") return kDOUBLE;
  SPRAY_FATAL("unknown data type " << s);
}

int PlyLoader::dataSize(int type) {
  switch (type) {
    case kCHAR: case kUCHAR: return 1;
    case kSHORT: case kUSHORT: return 2;
    case kINT: case kUINT: case kFLOAT: return 4;
    case kDOUBLE: return 8;
    default: SPRAY_FATAL("unknown data type " << type);
  }
}

void PlyLoader::quickHeaderRead(const std::string& filename, Header* header) {
  // the header is small: read a bounded prefix
  FILE* f = std::fopen(filename.c_str(), "rb");
  SPRAY_CHECK(f != nullptr, filename);
  std::vector<char> buf(1 << 16);
  const std::size_t size = std::fread(buf.data(), 1, buf.size(), f);
  std::fclose(f);

  std::size_t off = 0, num_vertices = 0, num_faces = 0;
  std::string line, word;
  SPRAY_CHECK(nextLine(buf.data(), size, &off, &line) && line == "ply", "unknown file type");
  int element_name = kUNDEFINED_ELEMENT_NAME, color_components = 0;
  while (nextLine(buf.data(), size, &off, &line)) {
    std::istringstream ss(line);
    word.clear();
    ss >> word;
    if (word == "element") {
      ss >> word;
      if (word == "vertex") { ss >> num_vertices; element_name = kVERTEX; }
      else if (word == "face") { ss >> num_faces; element_name = kFACE; }
      else element_name = kUNDEFINED_ELEMENT_NAME;
    } else if (word == "property") {
      if (element_name == kVERTEX) {
        std::string data_type, property_name;
        ss >> data_type >> property_name;
        if (property_name == "red" || property_name == "green" || property_name == "blue") ++color_components;
      }
    } else if (word == "end_header") {
      break;
    }
  }
  SPRAY_CHECK(num_vertices > 0, "");
  SPRAY_CHECK(num_faces > 0, "");
  header->num_vertices = num_vertices;
  header->num_faces = num_faces;
  header->has_color = (color_components == 3);
}

std::size_t PlyLoader::parseHeader(const char* buf, std::size_t size) {
  std::size_t off = 0;
  std::string line, word;
  SPRAY_CHECK(nextLine(buf, size, &off, &line) && line == "ply", "unknown file type");
  format_ = kUNDEFINED_FORMAT;
  num_vertices_ = num_faces_ = 0;
  elements_.clear();
  bool end_header = false;
  while (nextLine(buf, size, &off, &line)) {
    std::istringstream ss(line);
    word.clear();
    ss >> word;
    if (word == "format") {
      ss >> word;
      if (word == "ascii") format_ = kASCII;
      else if (word == "binary_little_endian") format_ = kLITTLE_ENDIAN;
      else if (word == "binary_big_endian") format_ = kBIG_ENDIAN;
      SPRAY_CHECK(format_ != kBIG_ENDIAN, "big endian PLY is untested in the reference and rejected");
    } else if (word == "element") {
      SPRAY_CHECK(format_ != kUNDEFINED_FORMAT, "element before format");
      SPRAY_CHECK(elements_.size() < 2, "more than 2 elements");  // MAX_NUM_ELEMENTS, ply_loader.cc:36
      Element e;
      std::string name;
      ss >> name >> e.num_elements;
      if (name == "vertex") { e.name = kVERTEX; num_vertices_ = e.num_elements; }
      else if (name == "face") { e.name = kFACE; num_faces_ = e.num_elements; }
      else SPRAY_FATAL("unknown element: " << name);
      elements_.push_back(e);
    } else if (word == "property") {
      SPRAY_CHECK(!elements_.empty(), "property before element");
      Element& e = elements_.back();
      if (e.name == kVERTEX) {
        std::string dtype, name;
        ss >> dtype >> name;
        if (name == "x" || name == "y" || name == "z") {
          if (e.vertex_dtype == kUNDEFINED_DATA_TYPE) {
            e.vertex_dtype = dataType(dtype);
            SPRAY_CHECK(e.vertex_dtype == kFLOAT, "vertex coordinates must be float");
          } else {
            SPRAY_CHECK(e.vertex_dtype == dataType(dtype), "inconsistent vertex types");
          }
          if (name == "x") e.last_property = P_X;
          else if (name == "y") { SPRAY_CHECK(e.last_property == P_X, "property order x,y,z"); e.last_property = P_Y; }
          else { SPRAY_CHECK(e.last_property == P_Y, "property order x,y,z"); e.last_property = P_Z; }
        } else if (name == "red" || name == "green" || name == "blue") {
          e.has_color = true;
          if (e.color_dtype == kUNDEFINED_DATA_TYPE) {
            e.color_dtype = dataType(dtype);
            SPRAY_CHECK(e.color_dtype == kUCHAR, "colours must be uchar");
          } else {
            SPRAY_CHECK(e.color_dtype == dataType(dtype), "inconsistent colour types");
          }
          if (name == "red") e.last_property = P_RED;
          else if (name == "green") { SPRAY_CHECK(e.last_property == P_RED, "property order red,green,blue"); e.last_property = P_GREEN; }
          else { SPRAY_CHECK(e.last_property == P_GREEN, "property order red,green,blue"); e.last_property = P_BLUE; }
        } else {
          SPRAY_FATAL("unknown property " << name);
        }
      } else {
        ss >> word;
        SPRAY_CHECK(word == "list", "face property must be a list");
        ss >> word; e.list_size_dtype = dataType(word);
        ss >> word; e.index_dtype = dataType(word);
        ss >> word;
        SPRAY_CHECK(word == "vertex_index" || word == "vertex_indices", "unknown property " << word);
      }
    } else if (word == "end_header") {
      end_header = true;
      break;
    }
  }
  SPRAY_CHECK(end_header, "no end_header");
  SPRAY_CHECK(format_ != kUNDEFINED_FORMAT, "no format");
  SPRAY_CHECK(num_vertices_ > 0, "");
  SPRAY_CHECK(num_faces_ > 0, "");
  return off;
}

void PlyLoader::load(const std::string& filename, Data* d) {
  const std::vector<char> file = slurp(filename);
  const char* buf = file.data();
  const std::size_t size = file.size();
  std::size_t off = parseHeader(buf, size);

  d->num_vertices = num_vertices_;
  d->num_faces = num_faces_;
  SPRAY_CHECK(d->vertices != nullptr, "");
  SPRAY_CHECK(d->faces != nullptr, "");

  for (const Element& e : elements_) {
    if (e.name == kVERTEX) {
      SPRAY_CHECK(e.num_elements * 3 <= d->vertices_capacity, "vertex capacity");
      if (e.has_color && d->colors) SPRAY_CHECK(e.num_elements <= d->colors_capacity, "colour capacity");
      if (format_ == kLITTLE_ENDIAN) {
        const std::size_t stride = 12 + (e.has_color ? 3 : 0);
        SPRAY_CHECK(off + stride * e.num_elements <= size, "truncated vertex data in " << filename);
        const unsigned char* p = (const unsigned char*)buf + off;
        if (!e.has_color) {
          std::memcpy(d->vertices, p, 12 * e.num_elements);  // x86 is little endian
        } else {
          for (std::size_t n = 0; n < e.num_elements; ++n, p += stride) {
            std::memcpy(d->vertices + 3 * n, p, 12);
            if (d->colors) d->colors[n] = ((uint32_t)p[12] << 16) | ((uint32_t)p[13] << 8) | (uint32_t)p[14];
          }
        }
        off += stride * e.num_elements;
      } else {  // ascii: one vertex per line
        std::string line;
        for (std::size_t n = 0; n < e.num_elements; ++n) {
          float x = 0.f, y = 0.f, z = 0.f;
          uint32_t r = 0, g = 0, b = 0;
          nextLine(buf, size, &off, &line);
          std::istringstream ss(line);
          ss >> x >> y >> z;
          d->vertices[3 * n] = x; d->vertices[3 * n + 1] = y; d->vertices[3 * n + 2] = z;
          if (e.has_color) {
            ss >> r >> g >> b;
            if (d->colors) d->colors[n] = (r << 16) | (g << 8) | b;
          }
        }
      }
    } else if (e.name == kFACE) {
      SPRAY_CHECK(e.num_elements * 3 <= d->faces_capacity, "face capacity");
      if (format_ == kLITTLE_ENDIAN) {
        const int cb = dataSize(e.list_size_dtype), ib = dataSize(e.index_dtype);
        SPRAY_CHECK(cb <= 4 && ib <= 4, "list types wider than 4 bytes");
        const std::size_t stride = (std::size_t)cb + 3 * (std::size_t)ib;
        SPRAY_CHECK(off + stride * e.num_elements <= size, "truncated face data in " << filename);
        const unsigned char* p = (const unsigned char*)buf + off;
        for (std::size_t n = 0; n < e.num_elements; ++n, p += stride) {
          int num_indices = 0;
          std::memcpy(&num_indices, p, (std::size_t)cb);
          SPRAY_CHECK(num_indices == 3, "face " << n << " has " << num_indices << " indices");
          uint32_t abc[3] = {0, 0, 0};
          for (int k = 0; k < 3; ++k) std::memcpy(&abc[k], p + cb + k * ib, (std::size_t)ib);
          d->faces[3 * n] = abc[0]; d->faces[3 * n + 1] = abc[1]; d->faces[3 * n + 2] = abc[2];
        }
        off += stride * e.num_elements;
      } else {
        std::string line;
        for (std::size_t n = 0; n < e.num_elements; ++n) {
          int num_indices = 0;
          uint32_t a = 0, b = 0, c = 0;
          nextLine(buf, size, &off, &line);
          std::istringstream ss(line);
          ss >> num_indices;
          SPRAY_CHECK(num_indices == 3, "face " << n << " has " << num_indices << " indices");
          ss >> a >> b >> c;
          d->faces[3 * n] = a; d->faces[3 * n + 1] = b; d->faces[3 * n + 2] = c;
        }
      }
    } else {
      SPRAY_FATAL("unknown element name " << e.name);
    }
  }
}

}  // namespace spray

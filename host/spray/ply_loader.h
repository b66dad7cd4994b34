// ply_loader.h — PLY reader in the dialect SpRay accepts (reference src/io/ply_loader.h:33-148,
// ply_loader.cc:42-503): ascii or binary_little_endian (big endian is CHECK-rejected there too),
// elements `vertex` (float x,y,z [uchar red,green,blue], in that order) and `face`
// (list <size type> <index type> vertex_index|vertex_indices, exactly 3 indices per face).
// Same Header / Data structs and the same fatal-on-malformed behaviour; colours are packed
// r<<16 | g<<8 | b (ply_loader.cc:247).
//
// Unlike the reference (one ifstream::read per field, ply_loader.cc:213-232,288-303) the file is read
// with ONE bulk read and decoded from memory — SURVEY.md section 8 row f1.
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

namespace spray {

class PlyLoader {
 public:
  struct Header {
    std::size_t num_vertices;  // out
    std::size_t num_faces;     // out
    bool has_color;            // out
  };

  struct Data {
    std::size_t vertices_capacity;  // in (floats)
    std::size_t faces_capacity;     // in (indices)
    std::size_t colors_capacity;    // in

    float* vertices;           // in/out
    std::size_t num_vertices;  // out

    uint32_t* faces;        // in/out
    std::size_t num_faces;  // out

    uint32_t* colors;  // rgb, in/out; may be nullptr
  };

  enum DataFormat { kUNDEFINED_FORMAT, kASCII, kBIG_ENDIAN, kLITTLE_ENDIAN };
  enum ElementName { kUNDEFINED_ELEMENT_NAME, kVERTEX, kFACE };
  enum DataType { kUNDEFINED_DATA_TYPE, kCHAR, kUCHAR, kSHORT, kUSHORT, kINT, kUINT, kFLOAT, kDOUBLE };

  struct Element {
    int name = kUNDEFINED_ELEMENT_NAME;
    std::size_t num_elements = 0;
    bool has_color = false;
    int vertex_dtype = kUNDEFINED_DATA_TYPE;
    int color_dtype = kUNDEFINED_DATA_TYPE;
    int last_property = 0;  // ordering check x->y->z, red->green->blue
    int list_size_dtype = kUNDEFINED_DATA_TYPE;
    int index_dtype = kUNDEFINED_DATA_TYPE;
  };

  // header only (ply_loader.cc:42-112)
  static void quickHeaderRead(const std::string& filename, Header* header);

  // full load into caller-provided arrays (ply_loader.cc:179-204)
  void load(const std::string& filename, Data* d);

  const std::vector<Element>& getElements() const { return elements_; }
  int getFormat() const { return format_; }

 private:
  std::size_t parseHeader(const char* buf, std::size_t size);  // returns the offset of the body
  static int dataType(const std::string& s);
  static int dataSize(int type);

  int format_ = kUNDEFINED_FORMAT;
  std::size_t num_vertices_ = 0, num_faces_ = 0;
  std::vector<Element> elements_;
};

}  // namespace spray

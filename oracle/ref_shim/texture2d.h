// oracle/ref_shim/texture2d.h — TEST INFRASTRUCTURE. Not part of the product.
//
// Stand-in for easycppogl/texture2d.h (an OpenGL texture wrapper; no graphics API exists on
// this box) so that bvh_gpu/gpu_bvh_scene.cpp compiles unmodified: the same create / alloc /
// update calls (texture2d.h:71-96,147-153) land in a host array instead of glTexImage2D /
// glTexSubImage2D. Texels are 4-byte components, rows of `width` texels, row 0 first.
#pragma once
#include <cstdint>
#include <cstring>
#include <map>
#include <memory>
#include <vector>

typedef int GLint;
typedef int GLsizei;
typedef unsigned int GLuint;
typedef unsigned int GLenum;
typedef unsigned char GLubyte;
typedef float GLfloat;

#ifndef GL_RGB32F
#define GL_RGBA32F 0x8814
#define GL_RGB32F 0x8815
#define GL_R32I 0x8235
#define GL_RGB32I 0x8D83
#define GL_NEAREST 0x2600
#endif

namespace EZCOGL {

class Texture2D;
using SP_Texture2D = std::shared_ptr<Texture2D>;

class Texture2D {
    GLint internal_ = 0;
    GLsizei width_ = 0, height_ = 0;
    int comps_ = 0;
    std::vector<uint32_t> texels_;  // raw 4-byte components (float or int32 bit patterns)

public:
    explicit Texture2D(const std::vector<GLint>& = {}) {}
    static SP_Texture2D create(const std::vector<GLint>& params = {}) { return std::make_shared<Texture2D>(params); }
    Texture2D(const Texture2D&) = delete;

    static int components(GLint internal) {
        switch (internal) {
            case GL_RGBA32F: return 4;
            case GL_RGB32F: return 3;
            case GL_RGB32I: return 3;
            case GL_R32I: return 1;
        }
        return 0;
    }
    void alloc(GLsizei w, GLsizei h, GLint internal, const GLubyte* ptr = nullptr) {
        internal_ = internal;
        width_ = w;
        height_ = h;
        comps_ = components(internal);
        texels_.assign(size_t(w) * size_t(h) * comps_, 0u);
        if (ptr) std::memcpy(texels_.data(), ptr, texels_.size() * 4);
    }
    template <typename T>
    void update(GLint x, GLint y, GLint w, GLint h, const T* data) {
        const unsigned char* src = reinterpret_cast<const unsigned char*>(data);
        for (GLint r = 0; r < h; ++r) {
            if (y + r >= height_) break;  // GL would raise GL_INVALID_VALUE and ignore the call
            std::memcpy(&texels_[(size_t(y + r) * width_ + x) * comps_], src + size_t(r) * w * comps_ * 4, size_t(w) * comps_ * 4);
        }
    }
    GLsizei width() const { return width_; }
    GLsizei height() const { return height_; }
    int comps() const { return comps_; }
    GLint internal() const { return internal_; }
    const uint32_t* raw() const { return texels_.data(); }
};

}  // namespace EZCOGL

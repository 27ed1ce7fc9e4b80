// camera.h — the viewer camera of the reference, reduced to the two matrices the
// traced path consumes (uniforms invV / invPV, montecarlo.cpp:439-440).
//
// Follows easycppogl/camera.h:62-99 and camera.cpp:28-95: perspective camera,
// field of view 0.78 rad, focal distance = scene_radius / tan(fov/2), modelview
// = T(0,0,-focal) * frame * T(-pivot), znear/zfar bracketing the scene sphere;
// matrices are evaluated in double and cast to float exactly as
// Camera::get_projection_matrix / get_modelview_matrix do. RTViewer multiplies
// the modelview by rotateX(-80) (montecarlo.cpp:404-405).
#pragma once
#include <algorithm>
#include <cmath>

#include "mcpt_math.h"

namespace mcpt {

class Camera {
    double field_of_view_ = 0.78;
    double asp_ratio_ = 1.0;
    double scene_radius_ = 1.0;
    double focal_dist_ = 1.0;
    double pivot_[3] = {0, 0, 0};

public:
    void set_aspect_ratio(double a) { asp_ratio_ = a; }
    void set_scene_radius(double r) {
        scene_radius_ = r;
        focal_dist_ = scene_radius_ / std::tan(field_of_view_ / 2.0);
    }
    void set_field_of_view(double fov) {
        field_of_view_ = fov;
        focal_dist_ = scene_radius_ / std::tan(field_of_view_ / 2.0);
    }
    void set_pivot_point(double x, double y, double z) {
        pivot_[0] = x;
        pivot_[1] = y;
        pivot_[2] = z;
    }
    double focal_dist() const { return focal_dist_; }

    // camera.cpp:28-39,56-64 with frame_ = identity
    Mat4 get_projection_matrix() const {
        double d = focal_dist_;
        double znear = std::max(0.01, d - scene_radius_);
        double zfar = d + scene_radius_;
        double range_inv = 1.0 / (znear - zfar);
        double f = 1.0 / std::tan(field_of_view_ / 2.0);
        double fx = (asp_ratio_ > 1) ? f / asp_ratio_ : f;
        double fy = (asp_ratio_ > 1) ? f : f * asp_ratio_;
        Mat4 P;
        for (int i = 0; i < 16; ++i) P.m[i] = 0.0f;
        P(0, 0) = float(fx);
        P(1, 1) = float(fy);
        P(2, 2) = float((znear + zfar) * range_inv);
        P(2, 3) = float(2 * znear * zfar * range_inv);
        P(3, 2) = -1.0f;
        return P;
    }
    // camera.cpp:64-70 with frame_ = identity
    Mat4 get_modelview_matrix() const {
        Mat4 V;
        V(0, 3) = float(-pivot_[0]);
        V(1, 3) = float(-pivot_[1]);
        V(2, 3) = float(-focal_dist_ - pivot_[2]);
        return V;
    }
};

// The matrices RTViewer::draw_ogl uploads for an untouched camera
// (montecarlo.cpp:388-389: centre 0, radius 145; :404-405,439-440).
inline void default_camera(int width, int height, float inv_v[16], float inv_pv[16]) {
    Camera cam;
    cam.set_scene_radius(145.0);
    cam.set_aspect_ratio(double(width) / double(height));
    Mat4 proj = cam.get_projection_matrix();
    Mat4 view = cam.get_modelview_matrix() * Transfo::rotateX(-80);
    Mat4 ipv = (proj * view).inverse();
    Mat4 iv = view.inverse();
    std::memcpy(inv_v, iv.data(), 16 * sizeof(float));
    std::memcpy(inv_pv, ipv.data(), 16 * sizeof(float));
}

}  // namespace mcpt

// host/camera.hxx -- raytracer::camera with the reference's constructor arguments (src/camera.hxx:24).
// The reference keeps its derived vectors private with no getters (src/camera.hxx:59-76); here the camera
// just records the arguments, and rt_camera_set derives lower_left_corner / horizontal / vertical on the
// host with the reference's float operation order.
#pragma once
#include "math.hxx"

namespace raytracer {
class camera final {
public:
    camera() = default;
    camera(math::vec3 position, math::vec3 lookat, math::vec3 up, float aspect, float vFOV, float aperture,
           float focus_distance) noexcept
        : position_{position}, lookat_{lookat}, up_{up}, aspect_{aspect}, vFOV_{vFOV}, aperture_{aperture},
          focus_distance_{focus_distance} {}

    math::vec3 const &position() const noexcept { return position_; }
    math::vec3 const &lookat() const noexcept { return lookat_; }
    math::vec3 const &up() const noexcept { return up_; }
    float aspect() const noexcept { return aspect_; }
    float vFOV() const noexcept { return vFOV_; }
    float aperture() const noexcept { return aperture_; }
    float focus_distance() const noexcept { return focus_distance_; }

    // camera::ray omits `- origin`, so it is only right with the eye at (0,0,0).  This returns the
    // same camera moved to the origin; translate the spheres by -position() alongside (scenes.hxx does).
    camera at_origin() const noexcept
    {
        return camera{math::vec3{0}, lookat_ - position_, up_, aspect_, vFOV_, aperture_, focus_distance_};
    }

private:
    math::vec3 position_{0}, lookat_{0, 0, -1}, up_{0, 1, 0};
    float aspect_{1.f}, vFOV_{72.f}, aperture_{0.f}, focus_distance_{1.f};
};
}  // namespace raytracer

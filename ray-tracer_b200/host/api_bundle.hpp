// The type bundle examples/scenes.hpp is instantiated with on the product side: the host mirror of the reference's scene API
// (scene_graph.hpp).  The oracle instantiates the same builders with its own f64 CPU types (oracle/oracle_capi.cpp).
#pragma once
#include "scene_graph.hpp"

namespace rtb {
namespace host {

struct HostApi {  // the type bundle scenes.hpp is instantiated with
    using Vec3 = rtb::host::Vec3;
    using Mat4 = rtb::host::Mat4;
    using BoundPtr = rtb::host::BoundPtr;
    using Sphere = rtb::host::Sphere;
    using Rectangle = rtb::host::Rectangle;
    using Cube = rtb::host::Cube;
    using Sprite = rtb::host::Sprite;
    using ConstantMedium = rtb::host::ConstantMedium;
    using BoundingVolumeHierarchyNode = rtb::host::BoundingVolumeHierarchyNode;
    using Lambertian = rtb::host::Lambertian;
    using Metal = rtb::host::Metal;
    using Dielectric = rtb::host::Dielectric;
    using DiffuseLight = rtb::host::DiffuseLight;
    using Isotropic = rtb::host::Isotropic;
    using ImageTexture = rtb::host::ImageTexture;
    static void setBakeSize(ImageTexture& t, uint32_t w, uint32_t h) { t.bakeWidth = w; t.bakeHeight = h; }
    using CheckerTexture = rtb::host::CheckerTexture;
    using Texture = rtb::host::Texture;
    using SolidColor = rtb::host::SolidColor;
    using HitList = rtb::host::HitList;
    using TransformedGeometry = rtb::host::TransformedGeometry;
};

}  // namespace host
}  // namespace rtb

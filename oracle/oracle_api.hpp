// ORACLE — TEST INFRASTRUCTURE ONLY.  API bundle that instantiates scenes/scenes.hpp with the oracle's types.
#pragma once
#include "rt_oracle.hpp"
#include "../scenes/scenes.hpp"

struct OracleApi {
    using Matrix = orc::Matrix;
    using Vector = orc::Vector;
    using Color = orc::Color;
    using Noise = orc::Noise;
    using Pattern = orc::Pattern;
    using Material = orc::Material;
    using ShapeArgs = orc::ShapeArgs;
    using Element = orc::Element;
    using GroupKind = orc::GroupKind;
    using PointLight = orc::PointLight;
    using World = orc::World;
    using Camera = orc::Camera;
    using ObjData = orc::ObjData;
    using ObjParser = orc::ObjParser;
};

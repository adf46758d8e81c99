// helios_b200/bench_components.hpp — the benchmark component set of SURVEY.md §8 (README.md:285-293,
// examples/ecs/query.cpp:16-29 in the reference), spelled so that TypeIndexOf<T>() == FNV-1a("helios::bench::<Name>"),
// the hashes the built-in system bodies (heliosengine_b200/csrc/builtin_systems.cu) and the Python harness use.
#pragma once
#include <cstdint>

namespace helios::bench {
struct Position { float x = 0, y = 0; };
struct Velocity { float dx = 0, dy = 0; };
struct Acceleration { float ddx = 0, ddy = 0; };
struct Mass { float m = 1; };
struct Health { int32_t hp = 0; };
struct Team { uint32_t id = 0; };
struct Damping { float k = 1; };
struct Frozen { uint32_t v = 0; };
struct Tint { uint32_t rgba = 0; };
struct CollisionEvent { uint64_t a = 0, b = 0; float impulse = 0; uint32_t pad = 0; };  // 24 B
}  // namespace helios::bench

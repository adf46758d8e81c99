// gpu_plugin_reference_headers.cpp — the GPU ECS plugin written against the REFERENCE'S OWN headers (not the shim in
// include/helios_b200/).  This is the variant an unmodified HeliosEngine host can `DynamicPlugin::Load`: its Build(App&)
// receives the reference's helios::app::App and uses only calls that class has.
//
// It is NOT built into libhelios_gpu_plugin.so in this image: the reference needs GCC 14+ / Clang 19+ (C++23 deducing-this)
// and Boost 1.87 / Taskflow 4.0 headers that are not installed.  tests/test_plugin_abi.py runs
//     g++ -std=c++23 -fsyntax-only -I/root/reference/src/<module>/include ... gpu_plugin_reference_headers.cpp
// as far as g++-13 gets and checks that no diagnostic points into THIS file; INTEGRATION.md §1 records the first error.
//
// Reference interfaces used (paths under /root/reference/src/):
//   app/include/helios/app/plugin.hpp:14-62                 Plugin virtuals
//   app/include/helios/app/dynamic_plugin.hpp:27,34-36,84-88 helios_create_plugin / helios_plugin_type_id, PluginTypeExport
//   app/include/helios/app/application.hpp:377-380          App::AddSystem(label, name, system)
//   app/include/helios/app/application.hpp:424,650          App::InsertResources, App::GetWorld
//   ecs/include/helios/ecs/system/param_traits.hpp:43-60    SystemParamTraits<P>::Make / RegisterAccess (custom parameters:
//                                                           ecs/README.md:106-205)
//   ecs/include/helios/ecs/world.hpp:491-620                World::InsertResources / WriteResource
#include <helios/app/application.hpp>
#include <helios/app/dynamic_plugin.hpp>
#include <helios/app/plugin.hpp>
#include <helios/app/schedules.hpp>
#include <helios/ecs/ecs.hpp>

#include <cstdint>
#include <string>
#include <utility>

#include "../../include/helios_gpu.h"

namespace helios::gpu {

// The benchmark components of configs[0] (README.md:285-312), spelled as a user of the reference would.
struct Position {
  float x = 0.0F;
  float y = 0.0F;
};
struct Velocity {
  float dx = 0.0F;
  float dy = 0.0F;
};

/// Resource that owns the device world and the schedule the GPU systems of one reference schedule live in.
/// Inserted with `app.InsertResources(GpuWorld{device})`; move-only like any owning resource.
struct GpuWorld {
  helios_gpu_world* world = nullptr;
  helios_gpu_schedule* update = nullptr;

  explicit GpuWorld(int device) {
    const helios_gpu_world_config cfg{device, 0, 0, HELIOS_GPU_WORLD_DEFAULT};
    if (helios_gpu_world_create(&cfg, &world) != HELIOS_GPU_OK) world = nullptr;  // no CUDA device: there is no CPU fallback
    if (world != nullptr) helios_gpu_schedule_create(world, "Update", &update);
  }
  GpuWorld(const GpuWorld&) = delete;
  GpuWorld& operator=(const GpuWorld&) = delete;
  GpuWorld(GpuWorld&& o) noexcept : world(std::exchange(o.world, nullptr)), update(std::exchange(o.update, nullptr)) {}
  GpuWorld& operator=(GpuWorld&& o) noexcept {
    std::swap(world, o.world);
    std::swap(update, o.update);
    return *this;
  }
  ~GpuWorld() {
    if (update != nullptr) helios_gpu_schedule_destroy(update);
    if (world != nullptr) helios_gpu_world_destroy(world);
  }
  [[nodiscard]] bool Valid() const noexcept { return world != nullptr; }
};

/// System parameter: a device query with the same component arguments as helios::ecs::Query<Args...>.
template <typename... Args>
struct GpuQuery {
  helios_gpu_world* world = nullptr;
  helios_gpu_schedule* schedule = nullptr;
};

}  // namespace helios::gpu

// Custom system parameter (ecs/README.md:106-205): the access policy is the one Query<Args...> would declare, so the
// reference's scheduler orders this system against CPU systems exactly as if it iterated on the CPU.
template <typename... Args>
struct helios::ecs::SystemParamTraits<helios::gpu::GpuQuery<Args...>> {
  using ParamType = helios::gpu::GpuQuery<Args...>;

  static ParamType Make(World& world, SystemLocalData& /*data*/, const AccessPolicy& /*policy*/) {
    auto& gw = world.WriteResource<helios::gpu::GpuWorld>();
    return ParamType{gw.world, gw.update};
  }

  static constexpr void RegisterAccess(AccessPolicyBuilder& builder) {
    SystemParamTraits<Query<Args...>>::RegisterAccess(builder);  // system/param_traits.hpp:55-60
    builder.template WriteResources<helios::gpu::GpuWorld>();
  }
};

namespace helios::gpu {

/// README.md:305-312 ChangePosition as a reference system whose body runs one frame of the device schedule.
struct ChangePositionOnGpu {
  void operator()(GpuQuery<Position&, const Velocity&> q) const {
    if (q.schedule != nullptr) helios_gpu_schedule_run_and_wait(q.schedule, HELIOS_GPU_EXEC_GRAPH);
  }
};

class GpuEcsPlugin final : public app::Plugin {
public:
  void Build(app::App& app) override {
    GpuWorld gw(0);
    ready_ = gw.Valid();
    if (!ready_) return;  // IsReady() stays false: App::WaitForPluginsReady reports it (application.cpp:209-229)
    // components are registered by type hash = TypeIndex (utils/type_info.hpp:308-311): the host compiler spells the names
    helios_gpu_register_component(gw.world, utils::TypeIndex::From<Position>().Hash(), sizeof(Position), alignof(Position), nullptr);
    helios_gpu_register_component(gw.world, utils::TypeIndex::From<Velocity>().Hash(), sizeof(Velocity), alignof(Velocity), nullptr);
    helios_gpu_query_desc qd{};
    qd.n_components = 2;
    qd.components[0] = utils::TypeIndex::From<Position>().Hash();
    qd.components[1] = utils::TypeIndex::From<Velocity>().Hash();
    qd.writes[0] = 1;  // Position& -> write set, const Velocity& -> read set (query/details/traits.hpp:73-97)
    helios_gpu_query* query = nullptr;
    helios_gpu_system_fn integrate = nullptr;
    helios_gpu_system_desc sd{};
    if (helios_gpu_query_create(gw.world, &qd, &query) != HELIOS_GPU_OK ||
        helios_gpu_builtin_system("integrate", &integrate) != HELIOS_GPU_OK) {
      ready_ = false;
      return;
    }
    sd.name = "ChangePosition";
    sd.query = query;
    sd.launch = integrate;
    helios_gpu_schedule_add_system(gw.update, &sd, nullptr);
    helios_gpu_schedule_build(gw.update);
    app.InsertResources(std::move(gw));                                               // application.hpp:424
    app.AddSystem(app::kUpdate, std::string("ChangePositionOnGpu"), ChangePositionOnGpu{});  // application.hpp:377-380
  }
  [[nodiscard]] bool IsReady(const app::App& /*app*/) const noexcept override { return ready_; }
  void Destroy(app::App& /*app*/) override { ready_ = false; }

private:
  bool ready_ = false;
};

}  // namespace helios::gpu

extern "C" {

// dynamic_plugin.hpp:27 — the host takes ownership (`plugin_.reset(raw_plugin)`, :583); nullptr = kCreateFailed
helios::app::Plugin* helios_create_plugin() { return new (std::nothrow) helios::gpu::GpuEcsPlugin(); }

// dynamic_plugin.hpp:34-36,84-88 — static storage: valid for the lifetime of the library image
const helios::app::PluginTypeExport* helios_plugin_type_id() {
  static const helios::app::PluginTypeExport kExport{helios::utils::TypeIndex::From<helios::gpu::GpuEcsPlugin>().Hash(),
                                                     "helios::gpu::GpuEcsPlugin"};
  return &kExport;
}

}  // extern "C"

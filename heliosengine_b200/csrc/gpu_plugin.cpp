// gpu_plugin.cpp — libhelios_gpu_plugin.so: a dynamic plugin in the reference's format (seam B1/B2).
// Exports exactly the two symbols DynamicPlugin::Load resolves (src/app/include/helios/app/dynamic_plugin.hpp:27,84-88,
// example: examples/app/dynamic_plugin/plugins/example_plugin.cpp:37-53).  Build(App&) registers the GPU movement
// system of BASELINE configs[0] (README.md:305-312 ChangePosition) on kUpdate; the body is sm_100a device code inside
// libhelios_gpu.so, reached through helios_gpu_builtin_system.
#include "../../include/helios_b200/app.hpp"
#include "../../include/helios_b200/bench_components.hpp"

namespace helios::gpu {

class GpuEcsPlugin final : public app::Plugin {
public:
  void Build(app::App& app) override {
    using namespace helios::bench;
    auto& world = app.GetWorld();  // the device world was created with the App: device init is synchronous here
    world.RegisterComponent<Position>();
    world.RegisterComponent<Velocity>();
    helios_gpu_system_fn integrate = nullptr;
    world.Check(helios_gpu_builtin_system("integrate", &integrate));
    app.AddSystem<ecs::Query<Position&, const Velocity&>>(app::kUpdate, "ChangePosition", integrate);
    ready_ = true;
  }
  bool IsReady(const app::App&) const noexcept override { return ready_; }
  void Destroy(app::App&) override { ready_ = false; }

private:
  bool ready_ = false;
};

}  // namespace helios::gpu

extern "C" {

helios::app::Plugin* helios_create_plugin() {
  // the host takes ownership (`plugin_.reset(raw_plugin)`, dynamic_plugin.hpp:583); nullptr = kCreateFailed
  try {
    return new helios::gpu::GpuEcsPlugin();
  } catch (...) {
    return nullptr;
  }
}

const helios::app::PluginTypeExport* helios_plugin_type_id() {
  static constexpr const char kName[] = "helios::gpu::GpuEcsPlugin";
  static const helios::app::PluginTypeExport kExport{helios::ecs::Fnv1a(kName), kName};  // static: valid for the library's lifetime
  return &kExport;
}

}  // extern "C"

// plugin_abi_app.cpp — the App object for plugin_abi_host.cpp.  libhelios_gpu_plugin.so is compiled against
// include/helios_b200/app.hpp, so the App it is handed must be THAT class: this translation unit is the
// only part of the ABI test that includes it, and it hands the host an opaque pointer.
#include "../../include/helios_b200/app.hpp"

extern "C" helios::app::App* abi_test_make_app() {
  try {
    return new helios::app::App();
  } catch (...) {
    return nullptr;  // no CUDA device: there is no CPU fallback
  }
}
extern "C" void abi_test_destroy_app(helios::app::App* app) { delete app; }
extern "C" uint64_t abi_test_systems_in_update(helios::app::App* app) {
  auto& schedule = app->GetMainSubApp().GetSchedule(helios::app::kUpdate);
  schedule.Build();  // Schedule::Build (schedule.cpp:117-158): the execution order lists every registered system
  return schedule.ExecutionOrder().size();
}

// plugin_abi_host.cpp — a host that knows NOTHING of this repository's headers: it only holds the reference's own
// declarations of helios::app::Plugin (src/app/include/helios/app/plugin.hpp:14-62) and PluginTypeExport
// (dynamic_plugin.hpp:34-36), restated below, and does what DynamicPlugin::Load / Unload do (dynamic_plugin.hpp:399-445,
// 541-586): dlopen, resolve helios_create_plugin / helios_plugin_type_id, take ownership of the Plugin*, drive its five
// virtuals through the BASE pointer, delete it before dlclose.  If the plugin's vtable were laid out differently from the
// reference's declaration, the calls below would land in the wrong slots and the observable protocol would break:
//   IsReady() == false before Build, true after Build (GPU init done), false again after Destroy.
//   plugin_abi_test nogpu <plugin.so>   everything except Build (Build needs a device world: no CPU fallback)
//   plugin_abi_test gpu   <plugin.so>   all five, Build with the App object plugin_abi_app.cpp provides
#include <dlfcn.h>

#include <cstdint>
#include <cstdio>
#include <cstring>

namespace helios::app {

class App;  // opaque here, exactly as in plugin.hpp:11

class Plugin {  // plugin.hpp:14-62: virtual order is the ABI
public:
  virtual ~Plugin() = default;
  virtual void Build(App& app) = 0;
  virtual void Finish(App& /*app*/) {}
  virtual void Destroy(App& /*app*/) {}
  virtual void Poll(App& /*app*/) {}
  [[nodiscard]] virtual bool IsReady(const App& /*app*/) const noexcept { return true; }
};

struct PluginTypeExport {  // dynamic_plugin.hpp:34-36
  uint64_t hash = 0;
  const char* qualified_name = nullptr;
};

}  // namespace helios::app

// provided by plugin_abi_app.cpp (the only translation unit that sees this repository's App)
extern "C" helios::app::App* abi_test_make_app();
extern "C" void abi_test_destroy_app(helios::app::App*);
extern "C" uint64_t abi_test_systems_in_update(helios::app::App*);

static int g_fail = 0, g_checks = 0;
#define CHECK(x)                                                             \
  do {                                                                       \
    ++g_checks;                                                              \
    if (!(x)) { ++g_fail; std::printf("FAIL %s:%d  %s\n", __FILE__, __LINE__, #x); } \
  } while (0)

static uint64_t Fnv1a(const char* s) {
  uint64_t h = 14695981039346656037ULL;
  for (; *s; ++s) h = (h ^ uint64_t(static_cast<unsigned char>(*s))) * 1099511628211ULL;
  return h;
}

int main(int argc, char** argv) {
  if (argc < 3) { std::printf("usage: plugin_abi_test gpu|nogpu <plugin.so>\n"); return 2; }
  const bool gpu = std::strcmp(argv[1], "gpu") == 0;
  void* lib = dlopen(argv[2], RTLD_NOW | RTLD_LOCAL);
  CHECK(lib != nullptr);
  if (!lib) { std::printf("%s\n", dlerror()); return 1; }
  using CreateFn = helios::app::Plugin* (*)();                        // dynamic_plugin.hpp:27 CreatePluginFn
  using TypeFn = const helios::app::PluginTypeExport* (*)();          // dynamic_plugin.hpp:84 PluginTypeExportFn
  auto create = reinterpret_cast<CreateFn>(dlsym(lib, "helios_create_plugin"));
  auto type_id = reinterpret_cast<TypeFn>(dlsym(lib, "helios_plugin_type_id"));
  CHECK(create != nullptr && type_id != nullptr);
  const helios::app::PluginTypeExport* te = type_id();
  CHECK(te != nullptr && te->qualified_name != nullptr && te->hash == Fnv1a(te->qualified_name));
  CHECK(type_id() == te);  // static storage: the same pointer for the lifetime of the image

  helios::app::Plugin* plugin = create();  // the host takes ownership (dynamic_plugin.hpp:583)
  CHECK(plugin != nullptr);
  if (plugin) {
    alignas(64) static unsigned char no_app[64];  // the four optional virtuals never touch their App argument
    helios::app::App* app = gpu ? abi_test_make_app() : reinterpret_cast<helios::app::App*>(no_app);
    CHECK(app != nullptr);
    CHECK(!plugin->IsReady(*app));            // slot 6: not built yet
    plugin->Poll(*app);                       // slot 5: default body
    CHECK(!plugin->IsReady(*app));
    if (gpu) {
      const uint64_t before = abi_test_systems_in_update(app);
      plugin->Build(*app);                    // slot 2: registers the GPU movement system on kUpdate
      CHECK(plugin->IsReady(*app));
      CHECK(abi_test_systems_in_update(app) == before + 1);
    }
    plugin->Finish(*app);                     // slot 3: default body
    CHECK(plugin->IsReady(*app) == gpu);
    plugin->Destroy(*app);                    // slot 4: overridden: back to "not ready"
    CHECK(!plugin->IsReady(*app));
    delete plugin;                            // slots 0/1: the plugin's own destructor, before dlclose (dynamic_plugin.hpp:437)
    if (gpu) abi_test_destroy_app(app);
  }
  CHECK(dlclose(lib) == 0);
  std::printf("plugin_abi_test %s: %d checks, %d failed\n", argv[1], g_checks, g_fail);
  return g_fail;
}

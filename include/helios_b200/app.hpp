// helios_b200/app.hpp — the slice of helios::app the GPU hot path plugs into (header-only, C++20).
//
// Mirrors (paths relative to /root/reference/src/app/): Plugin (include/helios/app/plugin.hpp:14-62), the dynamic
// plugin C ABI `helios_create_plugin` / `helios_plugin_type_id` + PluginTypeExport
// (include/helios/app/dynamic_plugin.hpp:27,34-36,84-88), schedule labels kPreUpdate/kUpdate/kPostUpdate
// (include/helios/app/schedules.hpp:44-146), App::AddPlugin/AddSystem/Initialize/Update/Run
// (include/helios/app/application.hpp:215-448, src/application.cpp:131-320) and SubApp (include/helios/app/sub_app.hpp).
// Only what drives the ECS hot path is present: one main world on one GPU, labelled schedules run in label order, then
// World::Update() once per frame (world.hpp:950-955).  Windowing, asset, logging and async-executor plumbing of the
// reference App are outside the hot path and are not mirrored.
#pragma once

#include <chrono>
#include <functional>
#include <map>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "ecs.hpp"

namespace helios::app {

class App;

// plugin.hpp:14-62 — same virtual interface, same order, so the vtable layout matches under the Itanium ABI.
class Plugin {
public:
  virtual ~Plugin() = default;
  virtual void Build(App& app) = 0;
  virtual void Finish(App& /*app*/) {}
  virtual void Destroy(App& /*app*/) {}
  virtual void Poll(App& /*app*/) {}
  [[nodiscard]] virtual bool IsReady(const App& /*app*/) const noexcept { return true; }
};

// dynamic_plugin.hpp:34-36
struct PluginTypeExport {
  uint64_t hash;
  const char* qualified_name;
};
inline constexpr const char* kDefaultCreateSymbol = "helios_create_plugin";      // dynamic_plugin.hpp:27
inline constexpr const char* kDefaultPluginIdSymbol = "helios_plugin_type_id";

// schedules.hpp:44-146 — the per-frame labels the hot path uses, in execution order.
enum class ScheduleLabel : int { kPreUpdate = 0, kUpdate = 1, kPostUpdate = 2 };
inline constexpr ScheduleLabel kPreUpdate = ScheduleLabel::kPreUpdate;
inline constexpr ScheduleLabel kUpdate = ScheduleLabel::kUpdate;
inline constexpr ScheduleLabel kPostUpdate = ScheduleLabel::kPostUpdate;

enum class AppExitCode : int { kSuccess = 0, kFailure = 1 };

// How a sub-app's update relates to the main frame (scheduler.cpp:204-240 ExtractSubApp / ClassifySubApp):
//   kBlocking     wait for the sub-app, extract, update — every frame (the default)
//   kOverlapping  kAllowOverlappingUpdates: the update is left running on the sub-world's stream while the main world
//                 starts its next frame; an extract stage that finds it still running is SKIPPED, at most
//                 kMaxOverlappingUpdates times in a row (0 = no limit), after which the main loop waits
//   kAsync        kAsync: extraction every frame; a new update pass starts whenever the previous one has finished
enum class SubAppMode : int { kBlocking = 0, kOverlapping = 1, kAsync = 2 };

// sub_app.hpp — a world of its own plus labelled schedules; the main app is the first sub-app.
class SubApp {
public:
  explicit SubApp(int device = 0, uint32_t entity_capacity = 0) : world_(device, entity_capacity) {}
  ecs::World& GetWorld() noexcept { return world_; }
  ecs::Schedule& GetSchedule(ScheduleLabel label) {
    auto it = schedules_.find(label);
    if (it == schedules_.end()) {
      static const char* names[] = {"PreUpdate", "Update", "PostUpdate"};
      it = schedules_.emplace(label, std::make_unique<ecs::Schedule>(world_, names[static_cast<int>(label)])).first;
      built_ = false;
    }
    return *it->second;
  }
  void Build() {
    for (auto& [l, s] : schedules_) s->Build();
    built_ = true;
  }
  bool Built() const noexcept { return built_; }
  // One frame of this sub-app: every schedule in label order (RunAndWait = run + ApplyDeferred), then World::Update.
  void Update() {
    if (pending_) FinishUpdate();
    for (auto& [l, s] : schedules_) s->RunAndWait();
    world_.Update();
    ++updates_;
  }
  // The asynchronous form of Update for overlapping / async sub-apps: every schedule but the last runs to completion
  // (ApplyDeferred needs the host), the LAST one is only enqueued (Schedule::Run) and keeps the GPU busy on this
  // sub-world's stream while the caller goes on; FinishUpdate applies its deferred data and ends the frame.
  void BeginUpdate() {
    if (pending_) FinishUpdate();
    ecs::Schedule* last = schedules_.empty() ? nullptr : schedules_.rbegin()->second.get();
    for (auto& [l, s] : schedules_)
      if (s.get() != last) s->RunAndWait();
    if (last) last->Run();
    pending_ = true;
  }
  void FinishUpdate() {
    if (!pending_) return;
    if (!schedules_.empty()) schedules_.rbegin()->second->ApplyDeferred();
    world_.Update();
    pending_ = false;
    ++updates_;
  }
  // sub_app.hpp IsUpdating / WaitUntilFullyIdle
  bool IsUpdating() {
    if (!pending_) return false;
    uint32_t idle = 0;
    world_.Check(helios_gpu_world_is_idle(world_.Handle(), &idle));
    return idle == 0;
  }
  void WaitUntilFullyIdle() { FinishUpdate(); }
  void SetMode(SubAppMode mode, size_t max_extraction_skips = 0) { mode_ = mode; max_skips_ = max_extraction_skips; }
  SubAppMode Mode() const noexcept { return mode_; }
  size_t MaxExtractionSkips() const noexcept { return max_skips_; }
  uint64_t UpdateCount() const noexcept { return updates_; }
  uint64_t ExtractCount() const noexcept { return extracts_; }
  uint64_t SkippedExtractCount() const noexcept { return skipped_; }
  void SetExecutor(ecs::ExecutorKind kind) {
    for (auto& [l, s] : schedules_) s->SetExecutor(kind);
    kind_ = kind;
  }
  // sub_app.hpp:148-152,425-433: the function Extract calls to copy data out of the main world before this sub-app's
  // update (the main world is passed mutable here only because device queries are created against a World&).
  using ExtractFn = std::function<void(ecs::World& main_world, ecs::World& sub_world)>;
  void SetExtractFunction(ExtractFn fn) { extract_fn_ = std::move(fn); }
  // sub_app.hpp:213-220,699-...
  void Extract(ecs::World& main_world) {
    if (extract_fn_) extract_fn_(main_world, world_);
    ++extracts_;
  }

private:
  friend class App;
  SubAppMode mode_ = SubAppMode::kBlocking;
  size_t max_skips_ = 0, consecutive_skips_ = 0;
  bool pending_ = false;
  uint64_t updates_ = 0, extracts_ = 0, skipped_ = 0;
  ExtractFn extract_fn_;
  bool built_ = false;
  ecs::World world_;  // declared before the schedules: they (and their queries) are destroyed before the world
  std::map<ScheduleLabel, std::unique_ptr<ecs::Schedule>> schedules_;
  ecs::ExecutorKind kind_ = ecs::ExecutorKind::kMultiThreaded;
};

class App {
public:
  explicit App(int device = 0, uint32_t entity_capacity = 0) : device_(device), main_(device, entity_capacity) {}
  ~App() {
    for (auto it = plugins_.rbegin(); it != plugins_.rend(); ++it) (*it)->Destroy(*this);  // application.cpp DestroyPlugins
  }

  ecs::World& GetWorld() noexcept { return main_.GetWorld(); }
  SubApp& GetMainSubApp() noexcept { return main_; }

  // application.hpp AddPlugin: the app takes ownership (dynamic_plugin.hpp:583 `plugin_.reset(raw_plugin)`).
  App& AddPlugin(std::unique_ptr<Plugin> plugin) {
    plugins_.push_back(std::move(plugin));
    return *this;
  }
  template <typename P, typename... A>
  App& AddPlugins(A&&... args) { return AddPlugin(std::make_unique<P>(std::forward<A>(args)...)); }

  // application.hpp:356-380 AddSystem(label, system) for GPU systems.
  template <typename Q>
  ecs::Schedule::SystemHandle AddSystem(ScheduleLabel label, const std::string& name, helios_gpu_system_fn launch,
                                        const ecs::GpuSystemOptions& opt = {}) {
    return main_.GetSchedule(label).AddGpu<Q>(name, launch, opt);
  }
  ecs::Schedule::SystemHandle AddSystem(ScheduleLabel label, const std::string& name, helios_gpu_system_fn launch,
                                        const ecs::GpuSystemOptions& opt = {}) {
    return main_.GetSchedule(label).AddGpu(name, launch, opt);
  }

  // application.cpp:131-320: Build every plugin, poll until all are ready (50 us -> 8 ms back-off, :24-25,209-229),
  // Finish, build the schedules.
  void Initialize() {
    if (initialized_) return;
    for (auto& p : plugins_) p->Build(*this);
    auto wait = std::chrono::microseconds(50);
    for (;;) {
      bool ready = true;
      for (auto& p : plugins_) {
        p->Poll(*this);
        ready = ready && p->IsReady(*this);
      }
      if (ready) break;
      std::this_thread::sleep_for(wait);
      if (wait < std::chrono::milliseconds(8)) wait *= 2;
    }
    for (auto& p : plugins_) p->Finish(*this);
    main_.Build();
    initialized_ = true;
  }
  // application.hpp AddSubApp: a world of its own, fed by its extract function after every main-world frame.
  SubApp& AddSubApp(const std::string& name, uint32_t entity_capacity = 0) {
    auto it = sub_apps_.find(name);
    if (it == sub_apps_.end()) it = sub_apps_.emplace(name, std::make_unique<SubApp>(device_, entity_capacity)).first;
    return *it->second;
  }
  SubApp& GetSubApp(const std::string& name) { return *sub_apps_.at(name); }
  // scheduler.cpp:60-151: main update, then the extract stage (every sub-app copies what it needs out of the main world),
  // then the sub-app updates.
  void Update() {
    Initialize();
    main_.Update();
    for (auto& [name, subp] : sub_apps_) {
      SubApp& sub = *subp;
      if (!sub.Built()) sub.Build();
      bool fresh = false;
      switch (sub.mode_) {  // scheduler.cpp:204-240 ExtractSubApp
        case SubAppMode::kBlocking:
          sub.WaitUntilFullyIdle();
          sub.Extract(main_.GetWorld());
          fresh = true;
          break;
        case SubAppMode::kAsync:
          if (!sub.IsUpdating()) sub.FinishUpdate();  // the loop's previous pass is over: close it, start the next below
          sub.Extract(main_.GetWorld());
          break;
        case SubAppMode::kOverlapping:
          if (sub.IsUpdating() && (sub.max_skips_ == 0 || sub.consecutive_skips_ < sub.max_skips_)) {
            ++sub.consecutive_skips_;
            ++sub.skipped_;
            break;
          }
          sub.WaitUntilFullyIdle();
          sub.Extract(main_.GetWorld());
          sub.consecutive_skips_ = 0;
          fresh = true;
          break;
      }
      // scheduler.cpp:153-172 LaunchSubAppUpdates: blocking sub-apps update now; overlapping ones start a pass after a
      // fresh extract and leave it running; async ones start a pass whenever none is in flight
      if (sub.mode_ == SubAppMode::kBlocking) sub.Update();
      else if (sub.mode_ == SubAppMode::kOverlapping && fresh) sub.BeginUpdate();
      else if (sub.mode_ == SubAppMode::kAsync && !sub.pending_) sub.BeginUpdate();
    }
    ++frame_;
  }
  // Scheduler::StopAsyncUpdateLoops / WaitUntilFullyIdle of every sub-app: call before reading sub-worlds on the host.
  void WaitForSubApps() {
    for (auto& [name, sub] : sub_apps_) sub->WaitUntilFullyIdle();
  }
  // Runs `frames` frames (the reference runs until an AppExit message; a frame budget is the headless equivalent).
  AppExitCode Run(uint64_t frames) {
    for (uint64_t i = 0; i < frames; ++i) Update();
    return AppExitCode::kSuccess;
  }
  uint64_t FrameCount() const noexcept { return frame_; }

private:
  int device_ = 0;
  SubApp main_;  // declared first: destroyed after the plugins
  std::map<std::string, std::unique_ptr<SubApp>> sub_apps_;
  std::vector<std::unique_ptr<Plugin>> plugins_;
  bool initialized_ = false;
  uint64_t frame_ = 0;
};

}  // namespace helios::app

// The two symbols a dynamic plugin library exports (dynamic_plugin.hpp:84-88).
extern "C" {
HELIOS_GPU_API helios::app::Plugin* helios_create_plugin();
HELIOS_GPU_API const helios::app::PluginTypeExport* helios_plugin_type_id();
}

// sspp_bindings.cpp — pybind11 module `_sspp`: the reference's Python surface over the B200 path.
//
// Mirrors /root/reference/src/sspp_bindings.cpp:14-68 name for name: classes Spline{3,6,7,9} and
// SamplingPathPlanner{3,6,7,9} with initializePath / evaluate / computeArcLength / findBestPath /
// get_ctrl_pts / plan(start, end, sigma, limits, sample_count=50, check_points=50, init_points=10)
// -> (bool, list[SplineD]); sampleWithNoise / checkCollision stay bound and, as in the reference
// (their std::default_random_engine& / mjData* arguments have no Python type), raise TypeError.
// Additive (no existing name changes): plan_device(...) -> dict with host arrays and raw device
// pointers of the same results, set_seed, set_verbose, Spline.knots, Spline.__call__.
// Host C++ only: links libsspp_cuda.so (the C ABI); no CUDA headers, no torch types.
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <mutex>

#include "sspp_planner.h"

namespace py = pybind11;

namespace {

template <int kDOF>
std::array<double, kDOF> to_point(const py::object& o, const char* what) {
  py::array_t<double, py::array::c_style | py::array::forcecast> a = py::array_t<double, py::array::c_style | py::array::forcecast>::ensure(o);
  if (!a || a.size() != kDOF) throw py::type_error(std::string(what) + ": expected " + std::to_string(kDOF) + " numbers");
  std::array<double, kDOF> p;
  for (int i = 0; i < kDOF; ++i) p[i] = a.data()[i];
  return p;
}
template <int kDOF>
py::array_t<double> from_point(const std::array<double, kDOF>& p) {
  py::array_t<double> a(kDOF);
  for (int i = 0; i < kDOF; ++i) a.mutable_data()[i] = p[i];
  return a;
}
// kDOF x K view on column-major control points, kept alive by `owner` (reference_internal semantics)
template <int kDOF>
py::array_t<double> ctrl_view(const sspp::ControlPoints<kDOF>& c, py::handle owner) {
  return py::array_t<double>({(py::ssize_t)kDOF, (py::ssize_t)c.cols}, {(py::ssize_t)sizeof(double), (py::ssize_t)(sizeof(double) * kDOF)},
                             c.data.data(), owner);
}

// The planner as Python sees it: plus a pool of Spline objects for plan()'s result list.  The reference returns thousands of
// fresh Spline objects per call (sspp_bindings.cpp:46-48); creating a pybind11 instance costs ~0.25 us, so 1 900 successes cost
// more than five kernel launches.  The pool keeps the objects of the previous result list and hands them out again ONLY when
// nothing but the pool still refers to them (reference count 1): whatever the caller kept - a list, one path, a numpy view of
// its control points - is left alone and replaced by a fresh object, so a returned Spline never changes under its holder.
template <int kDOF>
struct PooledPlanner : sspp::SamplingPathPlanner<kDOF> {
  using Base = sspp::SamplingPathPlanner<kDOF>;
  using Spline = typename Base::Spline;
  using Base::Base;
  struct Entry { py::object obj; Spline* ptr; int busy = 0; };   // busy: consecutive calls that found the object still referenced
  std::vector<Entry> pool;
  std::mutex mu;   // plan() / plan_device() run without the GIL: calls on ONE planner from several Python threads take turns here

  py::list to_list(const typename Base::PlannedBlock& result) {
    // `ok, paths = planner.plan(...)` in a loop still holds the previous list while the call runs, so its objects come free
    // one call later: an entry stays in the pool while busy for up to two calls (two sets alternate, nothing is allocated in
    // the steady state); an object referenced for longer belongs to its holder and leaves the pool
    size_t kept = 0;
    for (size_t i = 0; i < pool.size(); ++i) {
      Entry& e = pool[i];
      e.busy = (Py_REFCNT(e.obj.ptr()) == 1) ? 0 : e.busy + 1;
      if (e.busy <= 2) { if (kept != i) pool[kept] = std::move(e); ++kept; }
    }
    pool.resize(kept);
    const size_t n = (size_t)result.n_success;
    py::list out(n);
    size_t at = 0;
    for (size_t i = 0; i < n; ++i) {
      while (at < pool.size() && pool[at].busy) ++at;
      if (at == pool.size()) {
        py::object o = py::cast(Spline(), py::return_value_policy::move);
        Spline* p = o.cast<Spline*>();
        pool.push_back(Entry{std::move(o), p, 0});
      }
      Entry& e = pool[at++];
      e.ptr->rebind(result.block, (int)i);
      PyList_SET_ITEM(out.ptr(), (py::ssize_t)i, e.obj.inc_ref().ptr());
    }
    if (at < pool.size()) {                            // free objects beyond what this call used: keep a bounded reserve
      size_t spare = 0, k2 = at;
      for (size_t i = at; i < pool.size(); ++i)
        if (pool[i].busy || spare++ < 4096) { if (k2 != i) pool[k2] = std::move(pool[i]); ++k2; }
      pool.resize(k2);
    }
    return out;
  }
};

template <int kDOF>
void bind_SamplingPathPlanner(py::module& m, const std::string& class_name, const std::string& splineclass_name) {
  using Planner = PooledPlanner<kDOF>;
  using Spline = typename Planner::Spline;

  py::class_<Spline>(m, splineclass_name.c_str())
      .def(py::init<>())
      .def("ctrls", [](py::object self) { return ctrl_view<kDOF>(self.cast<const Spline&>().ctrls(), self); })
      .def("knots", [](const Spline& s) { return py::array_t<double>((py::ssize_t)s.knots().size(), s.knots().data()); })
      .def("__call__", [](const Spline& s, double u) { return from_point<kDOF>(s(u)); }, py::arg("u"));

  py::class_<Planner>(m, class_name.c_str())
      .def(py::init([](const std::string& xml_string, int device) { return new Planner(xml_string, device); }), py::arg("xml_string"),
           py::arg("device") = 0)   // `device` is additive: the CUDA ordinal this planner lives on (one process per GPU)
      .def("initializePath",
           [](Planner& self, const py::object& start, const py::object& end, Spline& init_spline, int num_points) {
             return self.initializePath(to_point<kDOF>(start, "start"), to_point<kDOF>(end, "end"), init_spline, num_points);
           },
           py::arg("start"), py::arg("end"), py::arg("init_spline"), py::arg("num_points") = 10)
      .def("evaluate", [](const Planner& self, double u) { return from_point<kDOF>(self.evaluate(u)); }, py::arg("u"))
      .def("evaluate", [](const Planner& self, const Spline& s, double u) { return from_point<kDOF>(self.evaluate(s, u)); },
           py::arg("spline"), py::arg("u"))
      // bound for name compatibility; uncallable from Python exactly as in the reference
      .def("sampleWithNoise",
           [](Planner&, const Spline&, double, const py::object&, const py::object&) -> Spline {
             throw py::type_error("sampleWithNoise(): incompatible function arguments: `generator` (std::default_random_engine&) "
                                  "has no Python type; candidates are sampled on the device by plan()");
           },
           py::arg("init_spline"), py::arg("sigma"), py::arg("limits"), py::arg("generator"))
      .def("checkCollision",
           [](Planner&, const Spline&, int, const py::object&) -> bool {
             throw py::type_error("checkCollision(): incompatible function arguments: `data` (mjData*) has no Python type; "
                                  "collision checks run on the device inside plan()");
           },
           py::arg("spline"), py::arg("num_samples"), py::arg("data"))
      .def("computeArcLength", &Planner::computeArcLength, py::arg("spline"), py::arg("check_points"))
      .def("findBestPath", &Planner::findBestPath, py::arg("successful_paths"), py::arg("best_spline"), py::arg("check_points") = 10)
      .def("get_ctrl_pts", [](py::object self) { return ctrl_view<kDOF>(self.cast<const Planner&>().get_ctrl_pts(), self); })
      .def("plan",
           [](Planner& self, const py::object& start, const py::object& end, double sigma, const py::object& limits, int sample_count,
              int check_points, int init_points) {
             const auto s = to_point<kDOF>(start, "start"), e = to_point<kDOF>(end, "end"), lim = to_point<kDOF>(limits, "limits");
             typename Planner::PlannedBlock result;
             {   // the reference holds the GIL through its OpenMP loop (no call_guard, sspp_bindings.cpp:43); the launch + the
                 // wait for the device need none of Python, so other threads run meanwhile
               py::gil_scoped_release release;
               std::lock_guard<std::mutex> turn(self.mu);
               result = self.plan_block(s, e, sigma, lim, sample_count, check_points, init_points);
             }
             return py::make_tuple(result.found, self.to_list(result));
           },
           py::arg("start"), py::arg("end"), py::arg("sigma"), py::arg("limits"), py::arg("sample_count") = 50,
           py::arg("check_points") = 50, py::arg("init_points") = 10)
      // ---- additive: device-buffer path ----
      .def("plan_device",
           [](Planner& self, const py::object& start, const py::object& end, double sigma, const py::object& limits, int sample_count,
              int check_points, int init_points, bool fetch_ctrl) {
             const auto s = to_point<kDOF>(start, "start"), e = to_point<kDOF>(end, "end"), lim = to_point<kDOF>(limits, "limits");
             sspp::PlanBuffers* Bp;
             std::unique_lock<std::mutex> turn(self.mu, std::defer_lock);   // held until the buffers have been read out below; taken without the GIL
             { py::gil_scoped_release release; turn.lock(); Bp = &self.plan_device(s, e, sigma, lim, sample_count, check_points, init_points); }
             sspp::PlanBuffers& B = *Bp;
             py::dict d;
             d["n_candidates"] = B.n_cand;
             d["n_success"] = B.n_success;
             d["best_idx"] = B.best_idx;
             d["best_cost"] = B.best_cost;
             d["arc_length"] = py::array_t<float>((py::ssize_t)B.L.size(), B.L.data());
             d["cost"] = py::array_t<float>((py::ssize_t)B.cost.size(), B.cost.data());
             d["status"] = py::array_t<uint8_t>((py::ssize_t)B.status.size(), B.status.data());
             // sampled control points: the collision-free candidates' rows arrived with the results; every candidate's rows
             // (failed ones too) cost one more device-to-host copy and come only when asked for
             d["ctrl_success"] = py::array_t<float>({(py::ssize_t)B.n_success, (py::ssize_t)B.n_free, (py::ssize_t)kDOF}, B.succ_rows);
             d["success_idx"] = py::array_t<int32_t>((py::ssize_t)B.n_success, B.succ_idx);
             if (fetch_ctrl) d["ctrl"] = py::array_t<float>({(py::ssize_t)B.n_cand, (py::ssize_t)B.n_free, (py::ssize_t)kDOF}, self.all_params().data());
             py::dict dev;   // raw device addresses (float32 / uint8 / uint64), for __cuda_array_interface__ wrappers
             dev["arc_length"] = (uintptr_t)B.view.L; dev["cost"] = (uintptr_t)B.view.cost; dev["status"] = (uintptr_t)B.view.status;
             dev["ctrl"] = (uintptr_t)B.view.params; dev["sorted_keys"] = (uintptr_t)B.view.keys; dev["stream"] = (uintptr_t)B.view.stream;
             dev["device"] = self.device();
             d["device_ptrs"] = dev;
             return d;
           },
           py::arg("start"), py::arg("end"), py::arg("sigma"), py::arg("limits"), py::arg("sample_count") = 50,
           py::arg("check_points") = 50, py::arg("init_points") = 10, py::arg("fetch_ctrl") = false)
      .def("set_seed", &Planner::set_seed, py::arg("seed"))
      .def("set_verbose", &Planner::set_verbose, py::arg("verbose"));
}

}  // namespace

PYBIND11_MODULE(_sspp, m) {
  m.doc() = "Pybind11 bindings for SamplingPathPlanner (B200 candidate-scoring path)";
  bind_SamplingPathPlanner<3>(m, "SamplingPathPlanner3", "Spline3");
  bind_SamplingPathPlanner<6>(m, "SamplingPathPlanner6", "Spline6");
  bind_SamplingPathPlanner<7>(m, "SamplingPathPlanner7", "Spline7");
  bind_SamplingPathPlanner<9>(m, "SamplingPathPlanner9", "Spline9");
  // raw-pointer capsules of the reference (src/sspp_bindings.cpp:63-68); no MuJoCo objects exist here,
  // the capsule just carries the address as it does there
  m.def("create_model_capsule", [](uint64_t model) { return py::capsule(reinterpret_cast<void*>(model), "mjModel"); });
  m.def("create_data_capsule", [](uint64_t data) { return py::capsule(reinterpret_cast<void*>(data), "mjData"); });
}

// tsp_bindings.cpp — pybind11 module `_tsp`: Python surface of the CURRENT tsp::TaskSpacePlanner adapter
// (/root/reference/include/sspp/tsp.h:7-109) over the B200 path.  The reference's own src/tsp_bindings.cpp is
// stale (it names SolverStatus::MaxIterationsExceeded, GradientStep fields, an XML-string constructor,
// initializePath, path_from_via_pt — none of which exist in tsp.h / tsp_types.h any more) and does not compile;
// this module binds what tsp.h actually offers, with the constructor taking the scene FILE NAME in place of mjModel*.
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include "sspp_tsp.h"

namespace py = pybind11;
using tsp::Point;

namespace {
Point to_point(const py::object& o, const char* what) {
  auto a = py::array_t<double, py::array::c_style | py::array::forcecast>::ensure(o);
  if (!a || a.size() != tsp::kDOF) throw py::type_error(std::string(what) + ": expected 4 numbers (x, y, z, yaw)");
  Point p;
  for (int i = 0; i < tsp::kDOF; ++i) p[i] = a.data()[i];
  return p;
}
py::array_t<double> from_point(const Point& p) { return py::array_t<double>(tsp::kDOF, p.data()); }
py::array_t<double> from_vias(const tsp::ViaSet& v) {
  py::array_t<double> a({(py::ssize_t)v.size(), (py::ssize_t)tsp::kDOF});
  for (size_t i = 0; i < v.size(); ++i)
    for (int d = 0; d < tsp::kDOF; ++d) a.mutable_at(i, d) = v[i][d];
  return a;
}
tsp::ViaSet to_vias(const py::object& o) {
  auto a = py::array_t<double, py::array::c_style | py::array::forcecast>::ensure(o);
  if (!a || a.size() % tsp::kDOF) throw py::type_error("via set: expected an array of shape [k, 4]");
  tsp::ViaSet v(a.size() / tsp::kDOF);
  for (size_t i = 0; i < v.size(); ++i)
    for (int d = 0; d < tsp::kDOF; ++d) v[i][d] = a.data()[i * tsp::kDOF + d];
  return v;
}
py::array_t<double> ctrl_array(const sspp::ControlPoints<tsp::kDOF>& c) {
  py::array_t<double> a({(py::ssize_t)tsp::kDOF, (py::ssize_t)c.cols});
  for (int i = 0; i < tsp::kDOF; ++i)
    for (int j = 0; j < c.cols; ++j) a.mutable_at(i, j) = c(i, j);
  return a;
}
}  // namespace

PYBIND11_MODULE(_tsp, m) {
  m.doc() = "Pybind11 bindings for tsp::TaskSpacePlanner (B200 candidate-scoring path)";
  py::enum_<tsp::SolverStatus>(m, "SolverStatus")
      .value("Converged", tsp::SolverStatus::Converged).value("Failed", tsp::SolverStatus::Failed)
      .value("BelowFloor", tsp::SolverStatus::BelowFloor).value("MaxIter", tsp::SolverStatus::MaxIter)
      .value("Unknown", tsp::SolverStatus::Unknown);
  m.def("SolverStatustoString", &tsp::SolverStatustoString);

  py::class_<tsp::Spline>(m, "Spline")
      .def("__call__", [](const tsp::Spline& s, double u) { return from_point(s(u)); }, py::arg("u"))
      .def("ctrls", [](const tsp::Spline& s) { return ctrl_array(s.ctrls()); })
      .def("knots", [](const tsp::Spline& s) { return py::array_t<double>((py::ssize_t)s.knots().size(), s.knots().data()); });

  py::class_<tsp::PathCandidate>(m, "PathCandidate")
      .def_property_readonly("via", [](const tsp::PathCandidate& c) { return from_vias(c.via); })
      .def_readonly("status", &tsp::PathCandidate::status)
      .def_readonly("L", &tsp::PathCandidate::L)
      .def_readonly("C_nf", &tsp::PathCandidate::C_nf)
      .def_readonly("C_wf", &tsp::PathCandidate::C_wf);

  py::class_<tsp::TaskSpacePlanner>(m, "TaskSpacePlanner")
      .def(py::init([](const std::string& xml_path, const std::string& body_name, double stddev_initial, double stddev_min, double stddev_max,
                       double stddev_increase_factor, double stddev_decay_factor, double elite_fraction, int sample_count, int check_points,
                       int gd_iterations, int init_points, double collision_weight, double z_min, const py::object& limits_min,
                       const py::object& limits_max, bool enable_gradient_descent, double sigma_floor, double var_ema_beta, double mean_lr,
                       double max_step_norm, double floor_margin, double floor_penalty_scale, bool faithful_quirks, int device) {
             std::shared_ptr<sspp::CudaModel> model;
             try { model = std::make_shared<sspp::CudaModel>(xml_path); }
             catch (const std::exception& e) { throw std::runtime_error(std::string("Failed to load MuJoCo model: ") + e.what()); }
             return new tsp::TaskSpacePlanner(model, body_name, stddev_initial, stddev_min, stddev_max, stddev_increase_factor,
                                              stddev_decay_factor, elite_fraction, sample_count, check_points, gd_iterations, init_points,
                                              collision_weight, z_min, limits_min.is_none() ? tsp::ones(-2.0) : to_point(limits_min, "limits_min"),
                                              limits_max.is_none() ? tsp::ones(2.0) : to_point(limits_max, "limits_max"), enable_gradient_descent,
                                              sigma_floor, var_ema_beta, mean_lr, max_step_norm, floor_margin, floor_penalty_scale,
                                              faithful_quirks, device);
           }),
           py::arg("model"), py::arg("body_name"), py::arg("stddev_initial") = 0.3, py::arg("stddev_min") = 0.01, py::arg("stddev_max") = 0.5,
           py::arg("stddev_increase_factor") = 1.5, py::arg("stddev_decay_factor") = 0.95, py::arg("elite_fraction") = 0.3,
           py::arg("sample_count") = 50, py::arg("check_points") = 50, py::arg("gd_iterations") = 0, py::arg("init_points") = 3,
           py::arg("collision_weight") = 1.0, py::arg("z_min") = 0.0, py::arg("limits_min") = py::none(), py::arg("limits_max") = py::none(),
           py::arg("enable_gradient_descent") = false, py::arg("sigma_floor") = 0.0, py::arg("var_ema_beta") = 0.2, py::arg("mean_lr") = 0.5,
           py::arg("max_step_norm") = 0.1, py::arg("floor_margin") = 0.01, py::arg("floor_penalty_scale") = 10.0,
           py::arg("faithful_quirks") = true, py::arg("device") = 0)
      .def("plan", [](tsp::TaskSpacePlanner& p, const py::object& s, const py::object& e, bool it) { return p.plan(to_point(s, "start"), to_point(e, "end"), it); },
           py::arg("start"), py::arg("end"), py::arg("iterate_flag") = false)
      .def("plan_iterations", [](tsp::TaskSpacePlanner& p, const py::object& s, const py::object& e, int n, bool it) {
             return p.plan_iterations(to_point(s, "start"), to_point(e, "end"), n, it); },
           py::arg("start"), py::arg("end"), py::arg("n_iters"), py::arg("iterate_flag") = false)
      .def("any_success", &tsp::TaskSpacePlanner::any_success)
      .def("get_succesful_path_candidates", &tsp::TaskSpacePlanner::get_succesful_path_candidates)
      .def("get_failed_path_candidates", &tsp::TaskSpacePlanner::get_failed_path_candidates)
      .def("get_sampled_via_sets", [](const tsp::TaskSpacePlanner& p) { py::list l; for (auto& v : p.get_sampled_via_sets()) l.append(from_vias(v)); return l; })
      .def("get_sampled_via_pts", [](const tsp::TaskSpacePlanner& p) { return from_vias(p.get_sampled_via_pts()); })
      .def("get_via_pts", [](const tsp::TaskSpacePlanner& p) { return from_vias(p.get_via_pts()); })
      .def("get_current_mean", [](const tsp::TaskSpacePlanner& p) { return from_point(p.get_current_mean()); })
      .def("get_current_stddev", [](const tsp::TaskSpacePlanner& p) { return from_point(p.get_current_stddev()); })
      .def("get_limits_min", [](const tsp::TaskSpacePlanner& p) { return from_point(p.get_limits_min()); })
      .def("get_limits_max", [](const tsp::TaskSpacePlanner& p) { return from_point(p.get_limits_max()); })
      .def("evaluate", [](const tsp::TaskSpacePlanner& p, double u) { return from_point(p.evaluate(u)); }, py::arg("u"))
      .def_static("evaluate_spline", [](double u, const tsp::Spline& s) { return from_point(s(u)); }, py::arg("u"), py::arg("spline"))
      .def("get_path_pts", [](const tsp::TaskSpacePlanner& p, int n) { return from_vias(p.get_path_pts(n)); }, py::arg("n") = 10)
      .def("get_ctrl_pts", [](const tsp::TaskSpacePlanner& p) { return ctrl_array(p.get_ctrl_pts()); })
      .def("get_knot_vector", [](const tsp::TaskSpacePlanner& p) { return py::array_t<double>((py::ssize_t)p.get_knot_vector().size(), p.get_knot_vector().data()); })
      .def("spline_from_via", [](const tsp::TaskSpacePlanner& p, const py::object& v) { return p.spline_from_via(to_point(v, "via")); })
      .def("spline_from_vias", [](const tsp::TaskSpacePlanner& p, const py::object& v) { return p.spline_from_vias(to_vias(v)); })
      .def("set_verbose", &tsp::TaskSpacePlanner::set_verbose)
      .def("set_seed", &tsp::TaskSpacePlanner::set_seed)
      .def("reset", &tsp::TaskSpacePlanner::reset);
}

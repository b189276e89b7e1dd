// sspp_model.h — scene ingestion for the scoring path: an MJCF subset is parsed without MuJoCo
// into a flat FP64 model (FlatModel, an "mjModel-lite"), which compile_model() specialises for
// one planner (which coordinates move) into the FP32 kernel-parameter block ssppm::DevModel.
//
// Replaces, for this path only: mj_loadXML at /root/reference/include/sspp.h:47 and
// src/main_icra_benchmark.cpp:146, name lookups of include/utility.h:34-68, and MuJoCo's
// collision filtering (contype/conaffinity, same-weld, parent-child, <exclude>).
#pragma once
#include <string>
#include <utility>
#include <vector>
#include "sspp_math.h"

namespace sspp {

enum MjJoint { MJ_NONE = -1, MJ_FREE = 0, MJ_BALL = 1, MJ_SLIDE = 2, MJ_HINGE = 3 };

struct FlatModel {
  int nbody = 0, ngeom = 0, nq = 0;
  std::vector<std::string> body_name, geom_name;
  std::vector<int> body_parent, body_weld, body_jtype, body_qadr;
  std::vector<double> body_pos, body_quat;   // 3, 4 per body (frame in parent)
  std::vector<double> jnt_axis, jnt_pos, jnt_range;  // 3, 3, 2 per body
  std::vector<double> qpos0;
  std::vector<int> geom_body, geom_type, geom_contype, geom_conaffinity;
  std::vector<double> geom_size, geom_pos, geom_quat, geom_margin;  // 3,3,4,1 per geom
  std::vector<int> exclude;   // body id pairs, 2 per entry

  int body_id(const std::string& name) const;
};

// Parse MJCF text.  Supported subset: <compiler angle eulerseq>, nested <default> classes with
// childclass, body trees with pos + quat/euler/axisangle/zaxis/xyaxes, <frame> wrappers, hinge /
// slide / free joints (several hinge or slide joints in one body become a chain of one-joint
// bodies sharing a weld id), primitive geoms incl. fromto, <contact><exclude>, <include> (resolved
// against base_dir).  Throws std::runtime_error with a message on malformed or unsupported input
// (collidable meshes / height fields / ellipsoids, ball joints, joint ref, explicit contact pairs).
FlatModel parse_mjcf(const std::string& xml_text, const std::string& base_dir = std::string());
FlatModel load_mjcf(const std::string& path);   // error text starts like the reference's: "Failed to load ..."

// true when MuJoCo's filters let geoms g1, g2 generate contacts
bool pair_active(const FlatModel& m, int g1, int g2);

enum PlanMode { PLAN_JOINT = 0,   // planner A: q = first `dof` scalar coordinates (hinge / slide joints)
                PLAN_TASK = 1 };  // planner B: (x,y,z,yaw) of one free body

struct CompileInfo {
  int n_static_pairs = 0;     // pairs between geoms that never move (folded into static_cost)
  int n_active_pairs = 0;     // moving-vs-static pairs
  int n_unreachable_pairs = 0; // moving-vs-static pairs dropped because the workspace bound proves them contact-free
  std::vector<int> mgeom_src, sgeom_src;   // FlatModel geom ids behind DevModel::mg / ::sg
};

// Specialise the flat model.  mode JOINT: dof planned coordinates; mode TASK: `body` is the free
// body that is planned.  Throws std::runtime_error when limits of DevModel are exceeded or the
// planned coordinates are not of a supported joint type.
struct CompileOptions {
  bool sdf = false;            // static scene will be a signed-distance grid: build sphere samples instead of pairs
  double sphere_ratio = 0.5;   // sample-sphere radius / smallest half extent (boxes), see sample_spheres()
  bool use_center_dist = true; // TASK contact cost: centre-distance rule or penetration depth
};
ssppm::DevModel compile_model(const FlatModel& m, PlanMode mode, int dof, const std::string& body,
                              CompileInfo* info = nullptr, const CompileOptions& opt = CompileOptions());

// sphere samples (centre in the geom frame, radius) covering a primitive from the inside
void sample_spheres(int type, const double* size, double ratio, std::vector<double>* out /* x,y,z,r per sphere */);

// FP64 helpers shared by the loader and the host planner (not the oracle)
void quat_to_mat(const double* q, double* R);   // (w,x,y,z) -> row-major 3x3, normalises

}  // namespace sspp

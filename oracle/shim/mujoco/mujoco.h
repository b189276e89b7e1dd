/* stand-in for <mujoco/mujoco.h> — TEST INFRASTRUCTURE (oracle/_ref).
 *
 * MuJoCo is a third-party dependency of the reference (CMakeLists.txt:30, requirements.txt:3 mujoco==3.2.5) that is
 * absent from /root/reference, this container and the GPU box.  This header declares the slice of its C API that the
 * reference's hot-path headers consume (include/sspp.h, Collision.h, utility.h, sspp/tsp_collision_world.h) with
 * MuJoCo's names, enum values and field meanings, so those headers compile unmodified.  The implementation
 * (oracle/ref_driver.cpp) is oracle O1's FP64 restatement of mj_kinematics + mj_collision: body tree with one joint per
 * body, geom frames, contype/conaffinity + weld + parent-child + <exclude> filtering, a contact kept iff dist < margin,
 * ONE contact per geom pair (MuJoCo's box-box manifolds emit several).  mj_forward here computes positions and
 * contacts only - what the reference reads (ncon, contact[].dist/geom1/geom2, geom_xpos) - not dynamics.
 * Models come from a flat array view (mjshim_make_model); mj_loadXML always fails with a message. */
#ifndef ORACLE_SHIM_MUJOCO_H
#define ORACLE_SHIM_MUJOCO_H
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef double mjtNum;
typedef enum mjtObj_ { mjOBJ_UNKNOWN = 0, mjOBJ_BODY = 1, mjOBJ_XBODY = 2, mjOBJ_JOINT = 3, mjOBJ_DOF = 4, mjOBJ_GEOM = 5 } mjtObj;
typedef enum mjtJoint_ { mjJNT_FREE = 0, mjJNT_BALL = 1, mjJNT_SLIDE = 2, mjJNT_HINGE = 3 } mjtJoint;

typedef struct mjContact_ {
  mjtNum dist;        /* signed distance between the geoms, negative = penetration */
  int geom1, geom2;
  int geom[2];
} mjContact;

typedef struct mjModel_ {
  int nq, nbody, njnt, ngeom;
  int* body_jntadr;   /* first joint of the body, -1: none */
  int* jnt_type;      /* mjtJoint */
  int* jnt_qposadr;
  int* jnt_bodyid;
  void* shim;         /* the flat model behind this view (oracle/ref_driver.cpp) */
} mjModel;

typedef struct mjData_ {
  mjtNum* qpos;       /* nq */
  mjtNum* xpos;       /* nbody x 3 */
  mjtNum* xquat;      /* nbody x 4 */
  mjtNum* geom_xpos;  /* ngeom x 3 */
  mjtNum* geom_xmat;  /* ngeom x 9 */
  int ncon;
  mjContact* contact;
  void* shim;
} mjData;

mjModel* mj_loadXML(const char* filename, const void* vfs, char* error, int error_sz);
void mj_deleteModel(mjModel* m);
mjData* mj_makeData(const mjModel* m);
mjData* mj_copyData(mjData* dest, const mjModel* m, const mjData* src);
void mj_deleteData(mjData* d);
void mj_forward(const mjModel* m, mjData* d);
int mj_name2id(const mjModel* m, int type, const char* name);
const char* mj_id2name(const mjModel* m, int type, int id);

#ifdef __cplusplus
}
#endif
#endif

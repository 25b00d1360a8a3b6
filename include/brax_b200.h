/*
 * brax_b200.h -- C ABI of the B200-native Brax v1 physics step.
 *
 * This is the drop-in boundary for ONE path of the reference
 * (Prasaya/brax, a fork of google/brax 0.1.1):
 *
 *     brax.physics.system.System.step(qp, act) -> (qp, info)
 *         reference: brax/physics/system.py:244-247
 *         -> _pbd_step      system.py:254-325   (dynamics_mode "pbd")
 *         -> _spring_step   system.py:342-377   (dynamics_mode "legacy_spring")
 *
 * The reference has no FFI of its own (it is pure Python/JAX); these entry
 * points are what a JAX FFI custom call / a ctypes stub / a torch wrapper binds
 * (see INTEGRATION.md).  Plain pointers and sizes only, no torch/JAX types.
 *
 * Layout at the boundary is exactly what JAX hands over for a batched step
 * (SURVEY.md 8(b)): C-contiguous float32, env-major:
 *     pos[B][N][3] rot[B][N][4] vel[B][N][3] ang[B][N][3] act[B][A]
 * quaternions are (w,x,y,z).  All outputs are written in full every call.
 *
 * There is NO CPU fallback behind these entry points: every bx_step* call
 * launches the sm_100a kernels, and returns an error if no CUDA device exists.
 */
#ifndef BRAX_B200_H_
#define BRAX_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BX_ABI_VERSION 14

/* dynamics_mode (base.py:179-201) */
#define BX_MODE_PBD 0
#define BX_MODE_LEGACY_SPRING 1

/* joint kinds.  pbd: joints.py:260 (Revolute), :322 (Spherical);
 * legacy_spring: spring_joints.py:114 (Revolute), :160 (Universal),
 * :226 (Spherical). */
#define BX_JOINT_REVOLUTE 0
#define BX_JOINT_SPHERICAL 1
#define BX_JOINT_SPRING_REVOLUTE 2
#define BX_JOINT_SPRING_UNIVERSAL 3
#define BX_JOINT_SPRING_SPHERICAL 4

/* actuator kinds: actuators.py:94 (Torque), :75 (Angle) */
#define BX_ACT_TORQUE 0
#define BX_ACT_ANGLE 1

/* force kinds: forces.py:26 (Thruster), :69 (Twister) */
#define BX_FORCE_THRUSTER 0
#define BX_FORCE_TWISTER 1

/* contact functions: colliders.py:744 capsule_plane, :805 capsule_capsule,
 * :667 box_plane.  A sphere is a one-end capsule (colliders.py:922-929). */
#define BX_PAIR_CAPSULE_PLANE 0
#define BX_PAIR_CAPSULE_CAPSULE 1
#define BX_PAIR_BOX_PLANE 2
/* colliders.py:822-849 capsule_mesh against geometry.TriangulatedBox (the ('capsule', 'box') entry of
 * colliders.get, :898-899; geometry.py:193-199, 462-540): one contact row per triangle of the box. */
#define BX_PAIR_CAPSULE_BOX 3
/* colliders.py:762-802 capsule_clippedplane (the ('capsule', 'clipped_plane') entry of colliders.get, :907-908;
 * geometry.ClippedPlane geometry.py:215-239): capsule ends against a finite rectangle of a plane whose normal is
 * oriented towards the capsule's body; a contact point in front of one of the four side planes is open. */
#define BX_PAIR_CAPSULE_CLIPPED_PLANE 4
/* colliders.py:699-742 box_heightmap (the ('box', 'heightMap') entry of colliders.get, :895; geometry.HeightMap
 * geometry.py:292-308): the 8 box corners against the triangle of the height grid under each corner. */
#define BX_PAIR_BOX_HEIGHTMAP 5

#define BX_MAX_BODIES 64
#define BX_MAX_JOINTS 64
#define BX_MAX_ACTUATORS 64
#define BX_MAX_FORCES 32
#define BX_MAX_PAIRS 64
#define BX_MAX_CONTACTS 128 /* K: contact rows per env */
#define BX_MAX_GROUPS 8

/* One rigid body: bodies.py:39-43 and integrators.py:38-44.
 * inv_inertia is the INVERSE diagonal inertia (bodies.py:40). */
typedef struct bx_body {
  float mass;
  float inv_inertia[3];
  float pos_mask[3]; /* 1 - frozen.position */
  float rot_mask[3]; /* 1 - frozen.rotation; quat_mask = [1, rot_mask] */
} bx_body;

/* One joint: joints.py:38-77 / spring_joints.py:42-87.  Matrices row-major:
 * axis_p[r*3+c] is component c of row r.  limit[d*2+{0,1}] = min,max (rad). */
typedef struct bx_joint {
  int32_t kind;
  int32_t body_p, body_c;
  int32_t group; /* index of the reference's joint group (summation order) */
  float off_p[3], off_c[3];
  float axis_p[9], axis_c[9];
  float limit[6];
  float angular_damping;
  float scale_pos, scale_ang;                      /* pbd */
  float stiffness, spring_damping, limit_strength; /* legacy_spring */
} bx_joint;

/* One actuator: actuators.py:31-46.  act_index[d] < 0 means "masked"
 * (act_mask = 0, actuators.py:46,64). */
typedef struct bx_actuator {
  int32_t kind;
  int32_t joint; /* index into joints[] */
  int32_t group;
  int32_t act_index[3];
  float strength;
} bx_actuator;

/* One body force: forces.py:29-41, 72-84. */
typedef struct bx_force {
  int32_t kind;
  int32_t body;
  int32_t act_index[3];
  float strength;
} bx_force;

/* One candidate collidable pair of a collider group (colliders.py:891-1023,
 * geometry.py:86-93,126-136,246-288).  Contact rows of the system are ordered
 * group-major, pair-major, contact-in-pair (system.py:36-43).
 *   capsule_plane : pts[c] = capsule end c in the body-a frame (n_contacts of
 *                   them), radius_a.
 *   box_plane     : pts[0..7] = corners in the body-a frame.
 *   capsule_capsule: pts[0] = end_a, pts[1] = pos_a, pts[2] = end_b,
 *                   pts[3] = pos_b (segment half-vector and centre offset in
 *                   each body frame), radius_a, radius_b. n_contacts = 1.
 *   capsule_box   : pts[0] = end_a, pts[1] = pos_a, radius_a; tri[f][v] = vertex v of triangle f
 *                   of the box in the body-b frame (BaseMesh.faces, geometry.py:146-156),
 *                   tri_normal[f] its normal. n_contacts = 12 (one row per triangle).
 *   capsule_clipped_plane: pts[0..n_contacts-1] = capsule ends, radius_a (as capsule_plane); the plane in the
 *                   body-b frame (geometry.py:222-239): tri[0][0] = normal, tri[0][1] = pos, tri[0][2] = x,
 *                   tri[1][0] = y, tri[1][1] = {halfsize_x, halfsize_y, 0}. One-way only.
 *   box_heightmap : pts[0..7] = box corners in the body-a frame (as box_plane); hm_height = the hm_size x hm_size
 *                   grid `HeightMap.height` row-major (HOST pointer, copied at bx_system_create), hm_cell_size =
 *                   size / (hm_size - 1).  The grid is indexed height[u, -v] as the reference does (:726);
 *                   negative indices wrap once, anything still outside is clamped (jax gather semantics; the
 *                   reference calls collisions outside the map 'not well defined', :705-706). One-way only. */
typedef struct bx_pair {
  int32_t kind;
  int32_t group;   /* collider group: averaging is per group (colliders.py:231) */
  int32_t two_way; /* 0: OneWayCollider (b static), 1: TwoWayCollider */
  int32_t body_a, body_b;
  int32_t n_contacts;
  float friction;   /* friction_a * friction_b   (colliders.py:317) */
  float elasticity; /* elasticity_a * elasticity_b */
  float radius_a, radius_b;
  float pts[8][3];
  float tri[12][3][3];    /* capsule_box; capsule_clipped_plane: the plane (see above) */
  float tri_normal[12][3];
  const float* hm_height; /* box_heightmap only (else NULL) */
  int32_t hm_size;
  float hm_cell_size;
} bx_pair;

/* Optional env epilogue ("next" row N1): everything the reference's Env.step computes
 * AFTER sys.step, evaluated by the same launch from the state still in shared memory.
 *   BX_ENV_ANT       brax/envs/ant.py:222-281       (Ant.step, _get_obs)
 *   BX_ENV_HUMANOID  brax/envs/humanoid.py:246-340  (Humanoid.step, _get_obs, _center_of_mass)
 * The fields are the envs' constructor arguments (ant.py:173-196, humanoid.py:200-221).
 * Body 0 is the torso; the centre of mass runs over bodies [0, N-1) (humanoid.py:332-334).
 * obs layout (floats per env, bx_system_obs_size):
 *   ant:      [pos0.z | pos0.xyz] rot0 joint_angle[D]  vel0 ang0 joint_vel[D]
 *             [clip(info.contact.vel)[N,3] clip(info.contact.ang)[N,3]]   (use_contact_forces)
 *   humanoid: [pos0.z | pos0.xyz] rot0 joint_angle[D]  vel0 ang0 joint_vel[D]
 *             com_inr[N-1,3,3] com_vel[N-1,3] com_ang[N-1,3] qfrc_actuator[sum of actuator_width]
 *   D = sum of joint_free_dofs.  metrics (bx_system_num_metrics) in the order of the
 *   reference's dict: ant.py:244-255 (10), humanoid.py:267-277 (9). */
#define BX_ENV_NONE 0
#define BX_ENV_ANT 1
#define BX_ENV_HUMANOID 2
/*   BX_ENV_FETCH     brax/envs/fetch.py:54-122  (Fetch.step, _get_obs)
 *   BX_ENV_GRASP     brax/envs/grasp.py:73-186  (Grasp.step incl. the action map and palm move
 *                    BEFORE sys.step :75-85, _get_obs); pbd only
 * obs: fetch  torso_fwd torso_up |target_local| target_local_dir pos_local[N,3] vel_local[N,3]
 *             contacts[N]                                                     (10 + 7N floats)
 *      grasp  |object_local| dir |target_local| dir pos_local[N,3] vel_local[N,3] hand_to_obj
 *             hand_vel heading |obj_to_target| dir obj_heading contacts[N]    (20 + 7N floats)
 * metrics: fetch.py:83-88 (5), grasp.py:123-128 (5).  The random target teleport after a hit
 * (fetch.py:88-93, grasp.py:131-136, JAX PRNG) stays with the caller: `hits` is metric 0. */
#define BX_ENV_FETCH 3
#define BX_ENV_GRASP 4
#define BX_MAX_ACT 64 /* act_size of systems with an action map (grasp) */
typedef struct bx_env_desc {
  uint32_t struct_size; /* = sizeof(bx_env_desc), checked */
  int32_t kind;
  float dt;             /* config.dt: velocity = (after - before) / dt (ant.py:226) */
  float forward_reward_weight; /* humanoid.py:253; ant: 1 */
  float ctrl_cost_weight;
  float contact_cost_weight;   /* ant.py:237-238 */
  float healthy_reward;
  float healthy_z_min, healthy_z_max;
  int32_t terminate_when_unhealthy;
  int32_t use_contact_forces;        /* ant.py:271-279 */
  int32_t exclude_current_positions; /* ..._from_observation */
  int32_t joint_free_dofs[BX_MAX_JOINTS];   /* dofs of joint i that Joint.angle_vel returns (joints.py:216-225) */
  int32_t actuator_width[BX_MAX_ACTUATORS]; /* act_index.shape[-1] of actuator i's group (humanoid.py:314-319) */
  /* fetch / grasp (fetch.py:28-34, grasp.py:32-54) */
  int32_t ref_body;    /* Torso / HandPalm: origin of the egocentric frame */
  int32_t target_body; /* Target */
  int32_t object_body; /* grasp: Object */
  int32_t hand_body;   /* grasp: HandThumbProximal */
  int32_t touch_bodies[4]; /* grasp.py:116 */
  float target_radius;
  float target_distance; /* _random_target: dist = radius + distance * u0 */
  float target_height;   /* z = height * u2 (grasp); <= 0: z = 1 (fetch.py:131) */
  float act_min[BX_MAX_ACT], act_range[BX_MAX_ACT]; /* grasp.py:43-54: act = min + range * (act + 1) / 2 */
} bx_env_desc;

/* Optional reset tables ("next" row N3): the static part of
 * System.default_qp(joint_angle=..., joint_velocity=...) (system.py:112-242), default_index 0.
 *   base_*  the QP after the config's default qps only (system.py:118-131)
 *   joints  in the depth order of system.py:142-152.  take[d] = slot of the flat joint-angle /
 *           joint-velocity vector that dof d reads, num_dof = none (0) (system.py:156-168);
 *           rot / ref = euler_to_quat of the joint's rotation / reference_rotation (:176-179)
 *   probes  of bodies.min_z (bodies.py:62-98): z = pos.z + rotate(point, rot).z - radius of
 *           sphere centres, capsule ends and box corners; every tree of bodies without a
 *           default qp is lowered by the minimum over its probes (and 0.0 if tree_zero: bodies
 *           without colliders, planes) (system.py:211-240).  body_tree[b] = tree of body b, -1 none. */
#define BX_MAX_PROBES 128
#define BX_MAX_TREES 8
typedef struct bx_reset_joint {
  int32_t body_p, body_c;
  int32_t take[3];
  float off_p[3], off_c[3];
  float rot[4], ref[4];
} bx_reset_joint;
typedef struct bx_reset_probe {
  int32_t body, tree;
  float point[3];
  float radius;
} bx_reset_probe;
typedef struct bx_reset_desc {
  uint32_t struct_size; /* = sizeof(bx_reset_desc), checked */
  int32_t num_dof;      /* D: length of the joint_angle / joint_velocity vectors */
  int32_t num_joints, num_probes, num_trees;
  float base_pos[BX_MAX_BODIES][3], base_rot[BX_MAX_BODIES][4];
  float base_vel[BX_MAX_BODIES][3], base_ang[BX_MAX_BODIES][3];
  bx_reset_joint joints[BX_MAX_JOINTS];
  bx_reset_probe probes[BX_MAX_PROBES];
  int32_t body_tree[BX_MAX_BODIES];
  int32_t tree_zero[BX_MAX_TREES];
} bx_reset_desc;

/* The static description of one system; everything System.__init__ derives
 * from the config (system.py:53-84).  Built on the host by
 * brax_b200.extract.describe(sys) from a constructed reference System. */
typedef struct bx_system_desc {
  uint32_t struct_size; /* = sizeof(bx_system_desc), checked */
  uint32_t abi_version; /* = BX_ABI_VERSION, checked */
  int32_t dynamics_mode;
  int32_t substeps; /* config.substeps; pbd runs substeps/2 pairs */
  int32_t num_bodies, num_joints, num_actuators, num_forces, num_pairs;
  int32_t num_collider_groups;
  int32_t act_size; /* A: length of the act vector */
  float dt_sub;     /* config.dt / config.substeps (integrators.py:45) */
  float gravity[3];
  float vel_damp; /* exp(velocity_damping * dt_sub) (integrators.py:89) */
  float ang_damp; /* exp(angular_damping * dt_sub)  (integrators.py:92) */
  float baumgarte_erp;      /* erp*substeps/dt (colliders.py:109) */
  float collide_scale;      /* colliders.py:112 */
  float velocity_threshold; /* |g|*h*4 (colliders.py:114) */
  const bx_body* bodies;
  const bx_joint* joints;
  const bx_actuator* actuators;
  const bx_force* forces;
  const bx_pair* pairs;
  const bx_env_desc* env; /* optional env epilogue, NULL = none */
  const bx_reset_desc* reset; /* optional reset tables (bx_default_qp), NULL = none */
} bx_system_desc;

typedef struct bx_system bx_system; /* opaque; owns device copies of tables */

/* Optional in-kernel policy of a rollout ("next" row N2: "T steps per launch with actions [T,B,A] or an
 * in-kernel MLP policy"): the acting loop of training/acting.py:53-74 -- actions = policy(obs),
 * state = env.step(state, actions) -- with the policy evaluated INSIDE the launch between the env steps, so a
 * whole unroll is one launch and neither obs nor actions cross the boundary between steps.
 * The network is the reference's policy MLP (training/networks.py:83-102: dense layers, the activation after
 * every layer but the last) on the normalised observation (obs - mean) / std
 * (training/acme/running_statistics.py `normalize`), its output the parameters of NormalTanhDistribution
 * (training/distribution.py:131-158): loc = out[:A], scale = softplus(out[A:2A]) + min_std;
 *   action = tanh(loc)                      (deterministic, `noise` NULL -- the distribution's mode)
 *   action = tanh(loc + scale * noise)      (`noise` = the caller's standard normal draws [T,B,A])
 * params: per layer l the weight matrix W_l[sizes[l]][sizes[l+1]] row-major followed by the bias
 * b_l[sizes[l+1]] (flax Dense `kernel`, `bias`), all layers concatenated, float32, device memory. */
#define BX_POLICY_MAX_LAYERS 8
#define BX_POLICY_RELU 0
#define BX_POLICY_SWISH 1
#define BX_POLICY_TANH 2
typedef struct bx_policy {
  uint32_t struct_size;  /* = sizeof(bx_policy), checked */
  int32_t num_layers;    /* dense layers including the output layer, 1..BX_POLICY_MAX_LAYERS */
  int32_t sizes[BX_POLICY_MAX_LAYERS + 1]; /* sizes[0] = obs size O; sizes[l+1] = width of layer l (<= 256);
                                              the last is 2A (loc, scale) or A (loc only) */
  int32_t activation;    /* BX_POLICY_* of the hidden layers */
  float min_std;         /* NormalTanhDistribution min_std (0.001) */
  const float* params;   /* device */
  const float* obs_mean; /* device [O] or NULL */
  const float* obs_std;  /* device [O] or NULL */
  const float* noise;    /* device [T,B,A] or NULL */
} bx_policy;

/* Device-side I/O of one batched step.  All pointers are DEVICE pointers,
 * float32, C-contiguous.  Shapes: N bodies, A act_size, K contact rows. */
typedef struct bx_step_io {
  /* in  */ const float *pos, *rot, *vel, *ang, *act; /* [B,N,3|4|3|3] [B,A] */
  /* out */ float *pos_o, *rot_o, *vel_o, *ang_o;
  /* Info (base.py:136-153).  Any of these may be NULL to skip the write. */
  float *contact_vel, *contact_ang;   /* info.contact  [B,N,3] x2 */
  float *joint_vel, *joint_ang;       /* info.joint    [B,N,3] x2 */
  float *actuator_vel, *actuator_ang; /* info.actuator [B,N,3] x2 */
  float *contact_pos;                 /* [B,K,3] */
  float *contact_normal;              /* [B,K,3] */
  float *contact_penetration;         /* [B,K]   */
  /* Optional epilogue (both NULL = skipped): Joint.angle_vel of the NEW state for
   * every joint (joints.py:197-226, what every env's _get_obs calls right after the
   * step, envs/ant.py:259), before the free_dofs selection: [B,J,3], slots beyond a
   * joint's dof are 0.  J = num_joints in descriptor order. */
  float *joint_angle, *joint_velocity;
  /* Optional env epilogue (systems created with desc.env; all NULL = skipped): the
   * reference's Env.step outputs, State.obs [B,O], State.reward [B], State.done [B],
   * State.metrics as [B,M].  With bx_rollout they are [T,B,...]: written every step. */
  float *obs, *reward, *done, *metrics;
  /* Optional INPUT of the fetch / grasp epilogues: three uniform [0,1) numbers per env (and per
   * step of a rollout: [T,B,3]) -- the draws of `_random_target` (fetch.py:124-134: distance,
   * angle; grasp.py:188-198: distance, angle, height).  With it the launch also teleports the
   * target of every env that scored a hit (fetch.py:88-93, grasp.py:131-136) in the QP it
   * returns; NULL leaves the targets where they are (the caller teleports). */
  const float* env_rand;
  /* Optional auto-reset of the env epilogue (device path; first_pos NULL = off): what
   * `envs.create(name, episode_length, auto_reset=True)` wraps around Env.step
   * (brax/envs/__init__.py:74-90): EpisodeWrapper.step (wrappers.py:105-120, action_repeat 1) and
   * AutoResetWrapper.step (wrappers.py:131-148), evaluated per env inside the launch:
   *   steps = done_in ? 0 : steps_in;  <Env.step>;  steps += 1;
   *   over = episode_length > 0 && steps >= episode_length;
   *   truncation = over ? 1 - done : 0;  done = over ? 1 : done;
   *   done ? (qp, obs) = (first_qp, first_obs)
   * The QP / obs / done the launch returns are the wrapped ones; in a rollout the chain
   * (done, steps) -> next step stays inside the kernel and truncation is [T,B]. */
  const float *first_pos, *first_rot, *first_vel, *first_ang; /* state.info['first_qp'] [B,N,3|4|3|3] */
  const float* first_obs; /* state.info['first_obs'] [B,O] */
  const float* done_in;   /* state.done of the state passed in [B] */
  const float* steps_in;  /* state.info['steps'] [B] (NULL: zeros) */
  float* steps_out;       /* [B] */
  float* truncation;      /* state.info['truncation'] [B] ([T,B] in a rollout) */
  /* Optional in-kernel policy (systems with desc.env; bx_step / bx_rollout / bx_state_rollout; NULL = the
   * actions come from `act`).  With it `act` is ignored (may be NULL): step t's action is computed in the
   * launch from the obs of step t-1 (first_obs for envs the auto-reset restarted) -- obs0 [B,O] = state.obs
   * for the first step -- and written to act_out [T,B,A] if given. */
  const bx_policy* policy; /* HOST pointer to the descriptor (its array pointers are device pointers) */
  const float* obs0;
  float* act_out;
  /* Optional per-env collider friction [B] (device; NULL = the descriptor's): domain randomisation as
   * experimental/tracing/randomizers.py:26-51 (friction_randomizer) + tracing/wrappers.py:27-88 do it --
   * EVERY collider of env e has material.friction = friction[e], so a pair's friction is
   * friction[e] * friction[e] (colliders.py:277,315,392,459,596).  bx_step / bx_rollout / bx_state_rollout;
   * the host-buffer entry points (bx_step_host*) refuse it. */
  const float* friction;
  /* Optional per-env joint sockets [B,J,3] (device; NULL = the descriptor's): Joint.parent_offset of every joint
   * (bx_joint.off_p) for env e -- the joint domain randomisation of experimental/tracing/randomizers.py:54-97
   * (joint_randomizer: parent_offset + offsets[e] on the selected joints, the config's value on the others).
   * Device path only, like friction. */
  const float* joint_offset;
  float episode_length;   /* EpisodeWrapper.episode_length; 0 = no episode limit */
  /* Non-zero (systems with desc.env, bx_step only): no physics -- the launch evaluates
   * info = System.info(qp) of the state passed in (system.py:327-340 / :378-390: the contact rows and
   * info.contact = the summed Collider.apply(qp), the spring-style impulse of the pose; the joint /
   * actuator parts, which no obs reads, stay zero) and the env epilogue on it: obs = _get_obs(qp, info)
   * as Env.reset builds it (envs/ant.py:203-204, humanoid.py:230-231, fetch.py:41-42, grasp.py:60-61).
   * The QP outputs are the inputs; reward / done / metrics are those of a zero-length step. */
  int32_t observe;
  /* EpisodeWrapper.action_repeat (wrappers.py:106-111; auto-reset launches only, 0 / 1 = none): every
   * wrapped step runs `action_repeat` env steps on the same action, sums their rewards and applies
   * the episode / reset logic once after the last; steps advance by action_repeat.  act and the
   * outputs stay [T,...]; env_rand, if given, is [T * action_repeat, B, 3] (one draw per env step). */
  int32_t action_repeat;
} bx_step_io;

/* Creates a system on CUDA device `device` (tables uploaded once).  Returns 0
 * on success; non-zero + bx_last_error() for anything outside coverage --
 * fail loudly, never fall back (SURVEY.md 8(b) "Errors"). */
int bx_system_create(const bx_system_desc* desc, int device, bx_system** out);
void bx_system_destroy(bx_system* sys);

/* Number of contact rows K and bodies N, act size A of a created system. */
int bx_system_num_contacts(const bx_system* sys);
int bx_system_num_bodies(const bx_system* sys);
int bx_system_act_size(const bx_system* sys);
/* Env epilogue sizes: O floats of obs per env, M metrics (0 without desc.env). */
int bx_system_obs_size(const bx_system* sys);
int bx_system_num_metrics(const bx_system* sys);

/* System.default_qp(joint_angle=..., joint_velocity=...) (system.py:112-242) for B independent
 * environments: forward kinematics down the joint tree from per-env joint angles / velocities
 * [B,D] (what every env's reset draws, envs/ant.py:198-204), then each free tree lowered onto the
 * xy plane.  ONE launch on `stream`, device pointers, outputs [B,N,3|4|3|3].  Systems created
 * with desc.reset only.  D = bx_system_num_joint_dof. */
int bx_system_num_joint_dof(const bx_system* sys);
int bx_default_qp(const bx_system* sys, int64_t num_envs, const float* joint_angle,
                  const float* joint_velocity, float* pos, float* rot, float* vel, float* ang,
                  void* stream);
/* The same with per-env joint sockets: joint_offset [B,J,3] (device; NULL = bx_default_qp) is Joint.parent_offset of
 * every joint of env e in the order of bx_system_desc.joints -- `default_qp` of the System each env builds under the
 * joint domain randomisation (experimental/tracing/randomizers.py:54-97; system.py:199-206 reads parent_offset). */
int bx_default_qp_offsets(const bx_system* sys, int64_t num_envs, const float* joint_angle,
                          const float* joint_velocity, const float* joint_offset, float* pos, float* rot,
                          float* vel, float* ang, void* stream);

/* One System.step for B independent environments: ONE kernel launch on
 * `stream` (a cudaStream_t passed as void*; NULL = default stream).  No host
 * sync, no allocation: CUDA-graph capturable.  Replaces
 * jit(vmap(System.step)) (system.py:244, envs/wrappers.py:70,80). */
/* In-place stepping is allowed: pos_o / rot_o / vel_o / ang_o may alias pos / rot / vel / ang (a tile reads
 * the whole state of its 32 environments into shared memory before it writes anything; tiles own
 * disjoint environments). */
int bx_step(const bx_system* sys, int64_t num_envs, const bx_step_io* io,
            void* stream);

/* The same step with HOST buffers (pageable or pinned): copies in, launches,
 * copies out, synchronises; the batch is pipelined in chunks over three streams
 * so upload, step and download overlap.  This is the call a numpy/ctypes binding of
 * System.step makes and what bench.py's e2e number times. */
int bx_step_host(bx_system* sys, int64_t num_envs, const bx_step_io* host_io);

/* Asynchronous form of bx_step_host for page-locked buffers: launches the zero-copy step on internal
 * stream `slot` (0..2) and returns; bx_host_wait(sys, slot) blocks until that slot's work is done and
 * its outputs are in host memory.  Independent groups of environments stepped through different slots
 * overlap their PCIe traffic and compute (a single host-side chain step -> step is latency bound).
 * The reference analogue is JAX's asynchronous dispatch of jit(vmap(System.step)) followed by
 * block_until_ready / device_get (brax/envs/to_torch.py:56-64).  Pageable buffers are refused. */
int bx_step_host_async(bx_system* sys, int64_t num_envs, const bx_step_io* host_io, int32_t slot);
int bx_host_wait(bx_system* sys, int32_t slot);

/* The host-buffer step for callers that may NOT allocate (code under CUDA graph capture
 * elsewhere in the process, tight serving loops, harnesses that forbid cudaMalloc).
 * NO COUNTERPART IN THE REFERENCE.
 * Same inputs, outputs and results as bx_step_host (host pointers, pageable or pinned;
 * H2D, ONE launch, D2H, synchronise `stream`), but
 *   - the device staging area is a file-scope `static __device__` array of
 *     BX_STATIC_SCRATCH_BYTES inside the library (every GPU that loads the library holds it
 *     until exit) instead of a cudaMalloc'ed buffer;
 *   - everything runs on the caller's `stream` (NULL = default stream); no stream, event or
 *     memory is created: the call makes no cudaMalloc / cudaFree / allocating library call;
 *   - a batch whose staging exceeds the array is REJECTED with an error that names the limit
 *     (ant with the full Info: about 21 300 environments; use bx_step_host beyond it).
 * All calls on one device share the array; it keeps the previous call's contents (nothing in
 * it needs to start at zero: inputs are overwritten before the launch, every requested output
 * is written in full).  Calls that overlap on the GPU race: one call at a time per device. */
#define BX_STATIC_SCRATCH_BYTES (32ll << 20)
int bx_step_host_static(const bx_system* sys, int64_t num_envs, const bx_step_io* host_io,
                        void* stream);
int64_t bx_static_scratch_bytes(void); /* = BX_STATIC_SCRATCH_BYTES of the loaded library */

/* PCIe bytes one bx_step_host call moves (outputs that are constants of the
 * dynamics mode are filled on the host, not transferred). */
int bx_host_transfer_bytes(const bx_system* sys, int64_t num_envs, int64_t* h2d,
                           int64_t* d2h);
/* Page-locked host memory for bx_step_host buffers (pageable works, slower). */
void* bx_host_alloc(size_t bytes);
void bx_host_free(void* p);

/* Shared-memory footprint of one 32-environment tile of this system, without a
 * device (sizing / diagnostics): bytes = state rows + tables; rows[8] = quad rows
 * of {bodies, act, slot lists + staging, contact rows, pre-velocities,
 * info.contact, info.actuator, info.joint}.  Two tiles share an SM when
 * 2 * (bytes + 1024) <= 233472. */
int bx_desc_tile_bytes(const bx_system_desc* desc, int64_t* bytes, int32_t* rows);

/* T consecutive steps in ONE launch: act is [T,B,A]; the state stays in shared memory between
 * the steps; the QP after the last step and the Info of the LAST step are written, the env
 * outputs (obs / reward / done / metrics / truncation) of EVERY step as [T,B,...].  Mirrors
 * lax.scan(env.step) over actions (training/acting.py:53-74). */
int bx_rollout(bx_system* sys, int64_t num_envs, int32_t num_steps,
               const bx_step_io* io, void* stream);

/* Opaque device-resident state of B environments ("next" row N2: persistent rollout state).  The handle
 * owns the QP of the batch on the system's device in a layout of the library's choosing; it is stepped
 * IN PLACE, so a training / acting loop (training/acting.py:53-74: lax.scan over env.step, the carry
 * living on the device) never moves the state through the boundary: actions in, env outputs out.
 *   bx_state_create   allocates the state for num_envs environments (contents undefined until set)
 *   bx_state_set/get  copy the QP in / out in the boundary layout [B,N,3|4|3|3] (device pointers, async
 *                     on `stream`)
 *   bx_state_rollout  num_steps >= 1 steps in ONE launch on the state: io->act is [T,B,A]; io's QP
 *                     input / output pointers are ignored (must be NULL); every other field of
 *                     bx_step_io means what it means for bx_rollout (Info of the last step, env outputs
 *                     of every step, auto-reset inputs, env_rand ...). */
typedef struct bx_state bx_state;
int bx_state_create(const bx_system* sys, int64_t num_envs, bx_state** out);
void bx_state_destroy(bx_state* state);
int64_t bx_state_num_envs(const bx_state* state);
int bx_state_set(bx_state* state, const float* pos, const float* rot, const float* vel, const float* ang,
                 void* stream);
int bx_state_get(const bx_state* state, float* pos, float* rot, float* vel, float* ang, void* stream);
int bx_state_rollout(bx_state* state, int32_t num_steps, const bx_step_io* io, void* stream);

/* Kernel launch configuration (tuning knobs; defaults are chosen per system).
 * threads_per_block: multiple of 32 in [32, 512], 0 = keep.
 * contact_section: 0 = keep, 1 = four barrier-separated contact phases,
 * 2 = fused per-body contact section (pbd, one-way plane colliders with at most
 * two contact rows per body; refused otherwise).  Returns 0 on success. */
int bx_system_configure(bx_system* sys, int32_t threads_per_block,
                        int32_t contact_section);

/* Picks the CTA size (warps per 32-env tile) by timing the candidates on the
 * caller's own device buffers for this batch size (about 16 launches + one
 * synchronise; outputs are overwritten, inputs untouched).  Optional: without
 * it a cost model chooses.  Like XLA's autotuning of a jitted computation. */
int bx_system_autotune(bx_system* sys, int64_t num_envs, const bx_step_io* io,
                       void* stream);
/* Threads per CTA currently configured. */
int bx_system_threads(const bx_system* sys);

/* Number of kernel launches issued through this library since load (for
 * bench.py's gpu_launches claim). */
int64_t bx_launch_count(void);

/* Thread-local message for the last non-zero return on this thread. */
const char* bx_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* BRAX_B200_H_ */

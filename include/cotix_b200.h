/*
 * cotix_b200.h — C-ABI of the B200-native batched rigid-body step engine for cotix.
 *
 * This is the drop-in boundary. The reference (the `cotix` Python/JAX package) has no
 * FFI of its own: its hot path is a set of pure Python callables on JAX pytrees. Each
 * entry point below replaces one of those callables for a whole batch and cites the
 * reference interface it stands in for (paths relative to the reference root).
 *
 * Conventions (all entry points):
 *   - extern "C", plain pointers and sizes; no torch / jax types.
 *   - return 0 on success, a negative COTIX_E* code on error; cotix_last_error()
 *     returns a thread-local message for the last failure.
 *   - "device" entry points only ENQUEUE work on the given stream (cudaStream_t passed as
 *     void*); they never synchronise, never allocate, and are CUDA-Graph capturable.
 *   - "host" entry points (suffix _host) take HOST buffers, copy in, run, copy out and
 *     synchronise; they are the reference-facing convenience path and what bench.py's
 *     `e2e` number times.
 *   - "no value" is NaN, never an error (reference: cotix/_contacts.py:19-24).
 *   - There is no CPU fallback anywhere in this library.
 */
#ifndef COTIX_B200_H
#define COTIX_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define COTIX_ABI_VERSION 2

/* error codes */
#define COTIX_OK 0
#define COTIX_EINVAL (-1)   /* bad argument (null pointer, bad size, bad enum)            */
#define COTIX_ECUDA (-2)    /* a CUDA runtime call failed; see cotix_last_error()          */
#define COTIX_ESHAPE (-3)   /* shape-type pair not in the reference dispatch table         */
#define COTIX_ENODEV (-4)   /* no usable sm_100 device                                     */

/* ------------------------------------------------------------------------------------ */
/* Shapes (reference: cotix/_convex_shapes.py Circle :10-47, AABB :50-133,               */
/* AbstractPolygon :136-194).  One record = 16 floats = 64 B, so a warp reads 2 KB of     */
/* contiguous records.                                                                    */
/*   rec[0] type   0 = Circle, 1 = AABB, 2 = Polygon (value stored as a float)            */
/*   rec[1] n      polygon vertex count 3..6 (float); ignored otherwise                   */
/*   rec[2] radius circle radius; ignored otherwise                                       */
/*   rec[3] pad                                                                           */
/*   rec[4..15]    Circle: position = rec[4..5];  AABB: lower = rec[4..5], upper =        */
/*                 rec[6..7];  Polygon: vertices k = rec[4+2k .. 5+2k], ALREADY in the    */
/*                 order Polygon.__init__ leaves them (order_clockwise,                   */
/*                 cotix/_geometry_utils.py:60-67) — use cotix_order_clockwise().         */
/* ------------------------------------------------------------------------------------ */
#define COTIX_SHAPE_CIRCLE 0
#define COTIX_SHAPE_AABB 1
#define COTIX_SHAPE_POLYGON 2
#define COTIX_SHAPE_REC_FLOATS 16
#define COTIX_MAX_POLY 6

typedef struct CotixNarrowParams {
  float init_dir[2];     /* GJK start direction = random_direction(PRNGKey(1)),          */
                         /* cotix/_collisions.py:287-298 (a global constant there)       */
  int32_t gjk_max_iters; /* 32, cotix/_collisions.py:101                                 */
  int32_t epa_iters_circle_poly; /* 128, cotix/_contacts.py:163                           */
} CotixNarrowParams;

/* status bits written per pair / per env */
#define COTIX_ST_GJK_CAP 0x1u      /* GJK hit gjk_max_iters                               */
#define COTIX_ST_EPA_CAP 0x2u      /* EPA used every allowed iteration                    */
#define COTIX_ST_NAN 0x4u          /* a NaN reached the dynamic state                     */
#define COTIX_ST_GROUND_NAN 0x8u   /* a non-finite impulse hit the static ground body     */
#define COTIX_ST_CCP_ERR 0x10u     /* circle_vs_aabb eqx.error_if condition               */
                                   /* (cotix/_contacts.py:105-107) was true               */
#define COTIX_ST_ANGLE_RANGE 0x20u /* |angle| >= 1e9: sin/cos defined as NaN              */

/*
 * Narrowphase sweep: one contact query per pair (a[k], b[k]).
 * Replaces cotix/_contacts.py: circle_vs_circle :30-58, aabb_vs_aabb :61-96,
 * circle_vs_aabb :99-154, circle_vs_polygon :157-202, aabb_vs_polygon :270-291,
 * polygon_vs_polygon :294-315 (which call GJK cotix/_collisions.py:20-112,277-310 and
 * EPA :115-273,313-329).  Dispatch follows the key order of `_contact_funcs`
 * (cotix/_colliders.py:21-35): type(a) <= type(b) in the order Circle < AABB < Polygon;
 * any other order makes the pair's result NaN and sets bit 0x80 in its status byte.
 *   exists[k]  1 iff the ContactInfo is not NaN (ContactInfo.isnan(), _contacts.py:26)
 *   pen[2k..]  penetration_vector (zeros when no contact, as ContactInfo.nan())
 *   cp[2k..]   contact_point (NaN, NaN when no contact)
 *   status[k]  COTIX_ST_* bits; may be NULL
 * All pointers are DEVICE pointers; a, b are [n][16] floats.
 */
int cotix_narrowphase(void* stream, int64_t n, const float* a, const float* b,
                      uint8_t* exists, float* pen, float* cp, uint8_t* status,
                      const CotixNarrowParams* params);

/*
 * The same sweep as a wavefront pipeline (classify -> GJK per type class -> EPA per type class over
 * stream-compacted queues in a caller-provided DEVICE workspace): identical results, several
 * times faster on mixed batches because every warp runs one code path.  `workspace` must be
 * 256-byte aligned and at least cotix_narrowphase_workspace_bytes(n) long; n < 2^32.
 */
int64_t cotix_narrowphase_workspace_bytes(int64_t n);
int cotix_narrowphase_ws(void* stream, int64_t n, const float* a, const float* b,
                         uint8_t* exists, float* pen, float* cp, uint8_t* status,
                         const CotixNarrowParams* params, void* workspace, int64_t workspace_bytes);

/*
 * The wavefront sweep WITHOUT any allocation, for callers that cannot call cudaMalloc (code under
 * CUDA-graph capture, tight serving loops, harnesses that forbid it).  No counterpart in the
 * reference.  Same arguments and bit-identical results
 * as cotix_narrowphase_ws, but the scratch is a file-scope `static __device__` array inside the
 * library (99 B per pair, about 13 MB; every GPU that loads the library holds it until exit), so
 * n is limited: n > COTIX_NARROWPHASE_STATIC_MAX_PAIRS fails with COTIX_EINVAL and a message that
 * names the limit.  The call issues no cudaMalloc / cudaFree / allocating routine; it only enqueues
 * on `stream` (the first call on a device also looks the array's address up with
 * cudaGetSymbolAddress: make that call outside a capture).  ALL calls on a device share the array
 * and it keeps the last call's contents: each call zeroes the queue counters it needs first, and
 * two calls that overlap on the GPU (different streams, no ordering between them) race — order
 * them on one stream or with events.
 */
#define COTIX_NARROWPHASE_STATIC_MAX_PAIRS 131072
int cotix_narrowphase_static(void* stream, int64_t n, const float* a, const float* b,
                             uint8_t* exists, float* pen, float* cp, uint8_t* status,
                             const CotixNarrowParams* params);

/* Same query with HOST buffers: copies in, runs, copies out, synchronises. */
int cotix_narrowphase_host(int device, int64_t n, const float* a, const float* b,
                           uint8_t* exists, float* pen, float* cp, uint8_t* status,
                           const CotixNarrowParams* params);

/*
 * order_clockwise (cotix/_geometry_utils.py:60-67) for a batch of polygons, in place on
 * DEVICE records (non-polygon records are left untouched).  This is what
 * Polygon.__init__ (cotix/_convex_shapes.py:143-144,197-229) does to its vertices.
 */
int cotix_order_clockwise(void* stream, int64_t n, float* recs);

/* ------------------------------------------------------------------------------------ */
/* LunarLander (reference: cotix/_lunar_lander.py, examples/test_viz.py:25-44).           */
/*                                                                                        */
/* State = struct-of-arrays over the env axis: one float/uint32 buffer `state[f*ld + e]`  */
/* for field f < COTIX_LUNAR_NF and env e < B, ld >= B.  Bodies follow the reference      */
/* order (cotix/_lunar_lander.py:143): 0 lander, 1 right_leg, 2 left_leg.  The ground     */
/* (body 3) is static — mass = inertia = inf, position = angle = 0 — so only its terrain  */
/* is stored.  Per body the 6 evolving DynamicBody leaves (cotix/_bodies.py:189-206).     */
/* ------------------------------------------------------------------------------------ */
#define COTIX_LUNAR_NBODY 3
#define COTIX_F_PX 0
#define COTIX_F_PY 1
#define COTIX_F_VX 2
#define COTIX_F_VY 3
#define COTIX_F_ANG 4
#define COTIX_F_ANGVEL 5
#define COTIX_F_BODY(b, f) ((b) * 6 + (f)) /* 0..17                                        */
#define COTIX_F_HEIGHT0 18 /* 18..25: terrain heights[0..7], _lunar_lander.py:110-114       */
#define COTIX_F_KNOT0 26   /* 26..29: random x-knots positions[1],[2],[5],[6], :116-125     */
#define COTIX_F_KEY0 30    /* uint32: PRNG key word 0                                       */
#define COTIX_F_KEY1 31    /* uint32: PRNG key word 1                                       */
#define COTIX_F_STEPS 32   /* uint32: steps taken in this episode                           */
#define COTIX_F_STATUS 33  /* uint32: sticky COTIX_ST_* bits of this episode                */
#define COTIX_F_RETURN 34  /* float : undiscounted return of this episode so far            */
#define COTIX_F_SHAPING 35 /* float : previous shaping potential (NaN = none yet)           */
#define COTIX_LUNAR_NF 36
#define COTIX_LUNAR_NOBS 8

#define COTIX_RNG_THREEFRY 0 /* jax.random default (threefry2x32), bit-exact key chain      */
#define COTIX_RNG_PHILOX 1   /* Philox4x32-10 counter mode, same distributions              */

/* done[] codes */
#define COTIX_DONE_NO 0
#define COTIX_DONE_CRASH 1    /* lander body touched the ground                             */
#define COTIX_DONE_OOB 2      /* left the arena                                             */
#define COTIX_DONE_LANDED 3   /* both legs on the ground and at rest                        */
#define COTIX_DONE_TIMEOUT 4  /* max_episode_steps reached                                  */
#define COTIX_DONE_NAN 5      /* non-finite state                                           */

typedef struct CotixLunarParams {
  /* ---- literals of the reference step; defaults reproduce it exactly --------------- */
  float dt;               /* 1e-2   examples/test_viz.py:26                               */
  float gravity_dv[2];    /* (0,-0.002) added to the lander velocity each step, :27-31    */
  float init_dir[2];      /* GJK start direction, cotix/_collisions.py:287-298            */
  float baumgarte;        /* 0.3    cotix/_collision_resolution.py:105                    */
  float baumgarte_dt;     /* 0.01   cotix/_collision_resolution.py:115                    */
  float coin_p;           /* 0.5    cotix/_colliders.py:220                               */
  float spring_kp;        /* 1.0    cotix/_lunar_lander.py:189                            */
  float spring_kv0;       /* 0.1                                                          */
  float spring_kv;        /* 0.05                                                         */
  float leg_damping;      /* 0.95   cotix/_lunar_lander.py:208,211                        */
  float body_props[4][4]; /* per body: mass, inertia, elasticity, friction (:78-141)      */
  float lander_local[6][2];   /* Polygon6 local vertices in stored order (:30)            */
  float leg_local[2][4][2];   /* [0] right_leg (body 1), [1] left_leg (body 2) (:32-72)   */
  float joint_local[4][2];    /* lander-side anchors: left j1, left j2, right j1, right j2 */
  float leg_joint2_local[2];  /* (0, 0.4)  (:158,173)                                      */
  float lander_start[3];      /* x, y, angle = 0, 5, 0.01 (:74,78-87)                      */
  float leg_offset[2][2];     /* [0] right (-1.2,-0.4), [1] left (+1.2,-0.4) (:75-76)      */
  float ground_floor;         /* -10  (:129,131)                                           */
  float terrain_x_outer;      /* 100  (:117,124)                                           */
  float pad_half_width;       /* 2    (:120-121)                                           */
  float pad_y;                /* -2   (:112-113)                                           */
  int32_t gjk_max_iters;      /* 32                                                        */
  int32_t rng_mode;           /* COTIX_RNG_*                                               */
  /* ---- new env logic the north_star asks for; the reference has none of it ---------- */
  int32_t substeps;           /* K physics sub-steps per call with the same action (>=1)   */
  int32_t auto_reset;         /* 1: a finished env is reset inside the same launch          */
  int32_t judge_enabled;      /* 0: reward = 0, done only on NaN/timeout (parity runs)      */
  int32_t max_episode_steps;  /* 0 = no limit                                               */
  int32_t reset_perturb;      /* 1: randomise the lander pose at reset                      */
  float main_engine_dv;       /* lander dv per step at full main throttle                   */
  float side_engine_dv;       /* lateral dv per step at full side throttle                  */
  float side_engine_dw;       /* angular-velocity change per step at full side throttle     */
  float x_limit, y_limit;     /* arena: |x| < x_limit, y < y_limit                          */
  float land_speed, land_angvel; /* "at rest" thresholds for COTIX_DONE_LANDED             */
  float reset_x[2], reset_y[2], reset_angle[2]; /* uniform ranges (reset_perturb)           */
  float reset_v, reset_w;     /* uniform half-ranges for velocity / angular velocity        */
  uint32_t philox_seed[2];    /* stream seed for COTIX_RNG_PHILOX                           */
  int32_t solver_loops;       /* SimpleConstraintSolver(loops) (cotix/_constraint_solvers.py:4-17):  */
                              /* sequential passes over the four joint constraints per step; 1 =     */
                              /* LunarLander.step as written (cotix/_lunar_lander.py:193-204)        */
} CotixLunarParams;

/* Fill *p with the reference literals (leaving the new env logic disabled). */
int cotix_lunar_default_params(CotixLunarParams* p);

/*
 * reset(key): replaces the LunarLander constructor cotix/_lunar_lander.py:29-143.
 * keys: DEVICE uint32 [B][2]; state: DEVICE [COTIX_LUNAR_NF][ld].
 */
int cotix_lunar_reset(void* stream, int64_t B, const uint32_t* keys, float* state,
                      int64_t ld, const CotixLunarParams* p);

/*
 * step(state, action): replaces the composed step f(env, key) of
 * examples/test_viz.py:25-44 = ExplicitEulerPhysics.step (cotix/_physics_solvers.py:16-33)
 * -> gravity -> RandomizedCollider.resolve (cotix/_colliders.py:74-351) -> key split ->
 * LunarLander.step (cotix/_lunar_lander.py:145-218), `substeps` times, plus the judge /
 * auto-reset logic (template: cotix/_envs.py:37-132).
 *   state_in / state_out  DEVICE [NF][ld]; may alias
 *   action   DEVICE [2][ld] (main, side) or NULL for zero action
 *   reward   DEVICE [B] float or NULL;  done DEVICE [B] uint8 or NULL
 *   obs      DEVICE [COTIX_LUNAR_NOBS][ld] float or NULL
 *            (x, y, vx, vy, angle, angvel, right-leg contact, left-leg contact)
 */
int cotix_lunar_step(void* stream, int64_t B, const float* state_in, float* state_out,
                     int64_t ld, const float* action, float* reward, uint8_t* done,
                     float* obs, const CotixLunarParams* p);

/*
 * The same step run as a WAVEFRONT PIPELINE (csrc/lunar_pipeline.cu): 1 memset per call + 6 small dense
 * kernels with stream-compacted work queues in a caller-provided DEVICE workspace, instead of
 * one fused kernel.  Results are identical bit for bit; it is several times faster for large,
 * contact-rich batches.  K = params->substeps sub-steps run the four step stages K times (an env
 * that finishes in a sub-step is frozen for the rest of the call, rewards add up) and the reset
 * stage once.  Requires B < 2^27.  `workspace` must be 256-byte aligned and at least
 * cotix_lunar_workspace_bytes(B) long; its contents need not be preserved between calls.
 * Like every device entry point it only enqueues on `stream` and is CUDA-Graph capturable.
 */
int64_t cotix_lunar_workspace_bytes(int64_t B);
int cotix_lunar_step_ws(void* stream, int64_t B, const float* state_in, float* state_out,
                        int64_t ld, const float* action, float* reward, uint8_t* done,
                        float* obs, const CotixLunarParams* p, void* workspace,
                        int64_t workspace_bytes);

/*
 * n_steps consecutive steps IN PLACE on `state` in one enqueue: step t uses the action
 * actions + (t % n_actions) * 2 * ld (DEVICE [n_actions][2][ld], or NULL), i.e. the loop
 *   for _ in range(n_steps): env = f(env, key)          examples/test_viz.py:46-48
 * as one call.  With a workspace (pipeline) the batch is cut into ranges of envs that run their
 * n_steps on their own streams WITHOUT joining between steps, so the latency-bound sparse stages of
 * one range execute under the dense stages of the others; the streams are forked from and joined
 * back into `stream`, so the call is stream-ordered and CUDA-Graph capturable (cotix_graph_*):
 * capture it once and a rollout of n_steps steps is ONE graph launch.  workspace NULL: up to
 * COTIX_ROLLOUT_MAX_ENVS envs (and substeps <= 1) run all n_steps inside ONE launch with the envs
 * resident in their thread block (csrc/lunar_rollout.cu: no launch gaps, no queue counters -- the
 * small-batch, latency-bound case); larger batches n_steps launches of the fused kernel.  Same results
 * bit for bit on every route.  reward / done / obs (or NULL) hold the last step's values.
 */
#define COTIX_ROLLOUT_MAX_ENVS 4736 /* 148 SMs x 32 envs per block: one block per SM */
int cotix_lunar_steps_ws(void* stream, int64_t B, float* state, int64_t ld, const float* actions,
                         int32_t n_actions, int32_t n_steps, float* reward, uint8_t* done, float* obs,
                         const CotixLunarParams* p, void* workspace, int64_t workspace_bytes);

/*
 * Launch configuration of the pipeline (process-wide; a negative value leaves a setting unchanged):
 *   split          ranges of envs the batch is cut into, 1..COTIX_MAX_CHUNKS; 0 = automatic = 1
 *                  (measured best for device-resident stepping).  Results do not depend on it.
 *   pdl            programmatic dependent launch between the stages of a range (default 1)
 *   priorities     sparse stages (narrow, contact, reset) at the highest launch priority, dense
 *                  stages at the lowest (default 1; only with split > 1)
 *   item_capacity  slots of every range's collision-item queue; 0 = 4 * envs + 4096.  An env whose
 *                  items do not fit evaluates its pairs inside the contact stage (same results); a
 *                  small value is how the tests force that branch.
 * cotix_lunar_pipeline_counters: queue lengths of the last step, summed over the ranges the
 * workspace is cut into for batch B under the current configuration: out[0..4] = collision items,
 * live contacts, envs with a candidate, auto-resets, envs that overflowed the item queue
 * (measurement aid; copies to the host and synchronises `stream`).
 */
int cotix_lunar_pipeline_config(int split, int pdl, int priorities, int64_t item_capacity);
int cotix_lunar_pipeline_counters(void* stream, int64_t B, const void* workspace, int64_t* out /*[5]*/);

/*
 * CUDA-Graph capture of any sequence of DEVICE entry points of this library (they only enqueue):
 *   cotix_graph_begin_capture(stream); cotix_lunar_steps_ws(stream, ...); ...;
 *   cotix_graph_end_capture(stream, &g);   then   cotix_graph_launch(g, stream)  per replay.
 * This is the single dispatch per rollout that jax.jit gives the reference's loop
 * (examples/test_viz.py:24,46-48).  Capture is thread-local; one open capture per thread.
 */
typedef struct CotixGraph CotixGraph;
int cotix_graph_begin_capture(void* stream);
int cotix_graph_end_capture(void* stream, CotixGraph** out);
int cotix_graph_launch(CotixGraph* g, void* stream);
int cotix_graph_info(const CotixGraph* g, int32_t* nodes, int64_t* kernel_launches);
int cotix_graph_destroy(CotixGraph* g);

/*
 * Per-stage timing of the pipeline (measurement aid; not thread-safe): while enabled, every
 * cotix_lunar_step_ws call runs as ONE range without programmatic dependent launch and records CUDA
 * events between its stages on the launching stream (serial stage times).
 * cotix_lunar_pipeline_stage_ms writes the average milliseconds of integrate, narrow, contact,
 * finish, reset and the whole step
 * since profiling was enabled, and returns the number of steps averaged.
 */
#define COTIX_PIPELINE_STAGES 6
int cotix_lunar_pipeline_profile(int enable);
int64_t cotix_lunar_pipeline_stage_ms(double* out_ms /*[COTIX_PIPELINE_STAGES]*/);

/*
 * Host-facing env handle: owns the device state of B envs on `device`, keeps it
 * resident, and exposes reset/step with HOST buffers (pinned or pageable).  One step =
 * H2D(action) -> kernels -> D2H(obs, reward, done), software-pipelined over ranges of envs
 * (cotix_lunar_env_set_chunks); the call returns when the host buffers are complete.
 */
typedef struct CotixLunarEnv CotixLunarEnv;
int cotix_lunar_env_create(CotixLunarEnv** out, int device, int64_t B,
                           const CotixLunarParams* p);
int cotix_lunar_env_destroy(CotixLunarEnv* env);
/* launch strategy of the handle's steps: AUTO = pipeline when B >= 16384 */
#define COTIX_MODE_AUTO 0
#define COTIX_MODE_FUSED 1
#define COTIX_MODE_PIPELINE 2
int cotix_lunar_env_set_mode(CotixLunarEnv* env, int mode);
/*
 * Host-step chunking: cotix_lunar_env_step_host splits the batch into `chunks` ranges of envs.
 * In pipeline mode the ranges are the pipeline's own (cotix_lunar_pipeline_config `split`): H2D(action)
 * of a range, its kernels and D2H(obs, reward, done) go on the range's stream, so PCIe traffic in both
 * directions overlaps the kernels of the other ranges.  In fused mode H2D of range c+1, the kernel of
 * range c and D2H of range c-1 run on three streams.  Envs are independent, so the results do not
 * depend on `chunks`.  0 = automatic (ranges of >= 256K envs, at most 4).  obs / reward / done of cotix_lunar_env_step_host may each be NULL: that output is not
 * copied to the host (obs stays readable on the device) — obs is 86 % of the D2H bytes of a step.
 */
#define COTIX_MAX_CHUNKS 16
int cotix_lunar_env_set_chunks(CotixLunarEnv* env, int chunks);
/*
 * Optional (default off): when every host buffer of a chunked step is page-locked, the whole
 * step (copies on three streams + kernels) is captured once per set of buffers and replayed as
 * ONE CUDA graph — one driver call per step instead of ~50; with pageable buffers the same work
 * is enqueued call by call.
 */
int cotix_lunar_env_set_host_graphs(CotixLunarEnv* env, int enable);
int cotix_lunar_env_reset_host(CotixLunarEnv* env, const uint32_t* keys /*[B][2]*/);
int cotix_lunar_env_step_host(CotixLunarEnv* env, const float* action /*[2][B]*/,
                              float* obs /*[8][B]*/, float* reward /*[B]*/,
                              uint8_t* done /*[B]*/);
/* copy the resident state out to / in from a HOST [NF][B] buffer */
int cotix_lunar_env_get_state_host(CotixLunarEnv* env, float* state);
int cotix_lunar_env_set_state_host(CotixLunarEnv* env, const float* state);
/* device pointer of the resident state ([NF][B], ld = B) and the handle's stream */
float* cotix_lunar_env_state_ptr(CotixLunarEnv* env);
void* cotix_lunar_env_stream(CotixLunarEnv* env);
/*
 * Roll `n_steps` steps with a fixed DEVICE action buffer (or NULL) as ONE replayed CUDA
 * graph of `n_steps` kernel nodes (captured on first use per n_steps).  Sum of rewards
 * per env is accumulated into ret_sum[B] (DEVICE, may be NULL).
 */
int cotix_lunar_env_rollout_graph(CotixLunarEnv* env, int32_t n_steps,
                                  const float* action_dev);

/* ------------------------------------------------------------------------------------ */
/* Differentiable rollout (K5, BASELINE config 5): hand-written adjoint of the step, the     */
/* replacement for jax.grad / jax.vjp of f(env, key) (examples/test_viz.py:25-44) composed   */
/* over T steps.  The reference differentiates by tracing its Python through jax; here the   */
/* backward pass is one CUDA kernel per step (csrc/lunar_adjoint.cu) that reloads the state  */
/* before the step, replays it with the forward kernels' own device functions (so every     */
/* discrete decision is reproduced) and walks it backwards.  Differentiated: the 18 dynamic  */
/* leaves (COTIX_LUNAR_NDYN rows: body b, leaf l -> row 6*b + l, leaves px, py, vx, vy,      */
/* angle, angular velocity, i.e. rows 0..17 of the state) and the action.  Held fixed, as   */
/* jax.grad holds them: broadphase / GJK outcomes, the EPA expansion sequence, argmax /      */
/* argsort selections, coin flips, jr.choice, early-outs, clip saturation; the terrain and   */
/* the PRNG key carry no cotangent.  Requires substeps == 1 and auto_reset == 0 (a reset     */
/* cuts the dependence on the past; roll episodes separately).                               */
#define COTIX_LUNAR_NDYN 18
/* per-env flag bits (OR-ed into `flags`; the caller zeroes the buffer) */
#define COTIX_ADJ_REPLAY_MISMATCH 0x1u /* the replayed step differs from state_next        */
#define COTIX_ADJ_NONFINITE 0x2u       /* non-finite step: cotangent dropped (set to zero) */
#define COTIX_ADJ_CONTACT_SHIFT 8      /* bits 8..10: resolve_collision calls in the step   */
/*
 * Vector-Jacobian product of ONE step.
 *   state_in    DEVICE [NF][ld]   the state before the step (what cotix_lunar_step read)
 *   state_next  DEVICE [NF][ld] or NULL: the state the forward step wrote; when given, the
 *               replay is compared with it bit for bit (COTIX_ADJ_REPLAY_MISMATCH)
 *   action      DEVICE [2][ld] or NULL (as in the forward call)
 *   g_out       DEVICE [NDYN][ld] cotangent of the next state's dynamic leaves
 *   g_in        DEVICE [NDYN][ld] receives the cotangent of state_in's; may alias g_out
 *   g_action    DEVICE [2][ld] or NULL
 *   flags       DEVICE uint32 [B] or NULL
 */
int cotix_lunar_step_bwd(void* stream, int64_t B, const float* state_in, const float* state_next,
                         int64_t ld, const float* action, const float* g_out, float* g_in,
                         float* g_action, uint32_t* flags, const CotixLunarParams* p);
/*
 * Forward rollout that keeps the tape: tape is DEVICE [T+1][NF][ld]; slot 0 holds the initial
 * state on entry, slot t+1 receives the state after step t.  actions DEVICE [T][2][ld] or NULL.
 * workspace / workspace_bytes as for cotix_lunar_step_ws, or NULL / 0 for the fused kernel.
 */
int cotix_lunar_rollout(void* stream, int64_t B, int32_t T, float* tape, int64_t ld,
                        const float* actions, const CotixLunarParams* p, void* workspace,
                        int64_t workspace_bytes);
/*
 * Backward pass over the tape of cotix_lunar_rollout: on entry g holds dL/d(state_T) (DEVICE
 * [NDYN][ld]), on return dL/d(state_0); g_actions DEVICE [T][2][ld] or NULL receives
 * dL/d(action_t).  Only enqueues on `stream`, no synchronisation, graph-capturable.
 *
 * workspace (DEVICE, 256-byte aligned, optional).  Two things about a step's VJP do not depend on
 * the cotangent and therefore not on the order of the reverse sweep — they depend only on the state
 * before the step, which is on the tape — so they are computed for MANY steps in one launch over
 * (step, env) items instead of inside T small, latency-bound launches:
 *  1. the DECISION TAPE (lunar_decide_kernel): the replay's narrowphase and coin flips; per
 *     env-step the <= 4 resolve_collision calls with their ContactInfo and final EPA edge
 *     (COTIX_ADJ_TAPE_WORDS words);
 *  2. the TRANSPOSED JACOBIANS of the env-steps that resolved a contact (lunar_jacobian_kernel):
 *     the VJP is linear in the cotangent, so for those env-steps — the ones with the expensive,
 *     divergent contact adjoint — the 20 x 18 matrix (g_state; g_action) = M g_next is formed in
 *     advance and the sweep spends a matrix-vector product on them (room for one env-step in eight;
 *     the rest fall back to the decision tape).
 * jacobians != 0: 1 + 2, the workspace holds cotix_lunar_adjoint_workspace_bytes(ld, steps) bytes;
 * jacobians == 0: 1 only, cotix_lunar_adjoint_decisions_bytes(ld, steps) bytes; steps = T covers
 * the whole rollout, a smaller workspace (>= 1 step) is used for groups of steps.  With
 * workspace == NULL every step replays its own decisions inside the sweep.  NULL and decisions-only
 * give identical bits; the matrix form sums the same terms in another order (rounding-level
 * differences).
 */
#define COTIX_ADJ_TAPE_WORDS 33
int64_t cotix_lunar_adjoint_workspace_bytes(int64_t ld, int32_t T);
int64_t cotix_lunar_adjoint_decisions_bytes(int64_t ld, int32_t T);
int cotix_lunar_rollout_bwd(void* stream, int64_t B, int32_t T, const float* tape, int64_t ld,
                            const float* actions, float* g, float* g_actions, uint32_t* flags,
                            const CotixLunarParams* p, void* workspace, int64_t workspace_bytes,
                            int32_t jacobians);

/* ------------------------------------------------------------------------------------ */
/* Headless rasteriser (K6): the replacement for the pygame window of cotix/_viz.py — there  */
/* is no graphics stack on the box, so the 800 x 600 "screen" is an RGB8 buffer in DEVICE    */
/* memory, [n_img][600][800][3] uint8 (row 0 = top, as pygame).  Primitives are painted in   */
/* list order onto a black image (PyPainter.next clears to black, :43-49).  Pixel rules: see */
/* csrc/raster_kernel.cu; agreement with SDL's own rasteriser is unpinned (no pygame here).  */
#define COTIX_IMG_W 800
#define COTIX_IMG_H 600
#define COTIX_PRIM_WORDS 16
#define COTIX_PRIM_LINE 0    /* pygame.draw.line, width 1      cotix/_viz.py:23-26           */
#define COTIX_PRIM_CIRCLE 1  /* pygame.draw.circle, filled     cotix/_viz.py:17-21           */
#define COTIX_PRIM_POLYGON 2 /* pygame.draw.polygon, filled, convex, 3..6 vertices  :28-31   */
#define COTIX_LUNAR_PRIMS 44 /* draw calls of LunarLander.draw, cotix/_lunar_lander.py:220-225 */
/*
 * Primitive record = int32[16]: type, colour 0x00RRGGBB, n (circle: radius in pixels; polygon:
 * vertex count), 0, then up to 6 (x, y) PIXEL coordinates (line: 2, circle: the centre).
 *   prims  DEVICE int32 [n_img][max_prims][16];  counts DEVICE int32 [n_img] or NULL (= max_prims)
 */
int cotix_raster_draw(void* stream, int64_t n_img, const int32_t* prims, const int32_t* counts,
                      int32_t max_prims, uint8_t* image);
/*
 * env.draw(painter) for LunarLander envs (cotix/_lunar_lander.py:220-225 through
 * Polygon.drawEdges cotix/_convex_shapes.py:189-194 and PyPainter._scale_vector
 * cotix/_viz.py:33-38): image i shows env env_index[i] (DEVICE int64 [n_img], or NULL for envs
 * 0..n_img-1) of `state` (DEVICE [NF][ld]).  cotix_lunar_prims writes the same 44 draw calls as
 * primitive records (DEVICE int32 [n_img][44][16]) instead of painting them.
 */
int cotix_lunar_draw(void* stream, int64_t n_img, const float* state, int64_t ld,
                     const int64_t* env_index, uint8_t* image, const CotixLunarParams* p);
int cotix_lunar_prims(void* stream, int64_t n_img, const float* state, int64_t ld,
                      const int64_t* env_index, int32_t* prims, const CotixLunarParams* p);

/* ------------------------------------------------------------------------------------ */
/* Generic scene step (K2): bodies with circle / AABB / polygon parts, e.g. the RoboCup     */
/* scene of cotix/_robocup.py:14-130.  One call = ExplicitEulerPhysics.step                 */
/* (cotix/_physics_solvers.py:16-33) + RandomizedCollider.resolve                           */
/* (cotix/_colliders.py:74-351) (+ key, _ = split(key), examples/test_viz.py:39) for B envs */
/* that share the static description and differ in their dynamic state and key.             */
/* ------------------------------------------------------------------------------------ */
#define COTIX_SCENE_MAX_BODIES 20
#define COTIX_SCENE_MAX_PARTS 28
#define COTIX_SCENE_MAX_PAIRS 254    /* distinct (part, part) evaluations with body_i >= body_j */
#define COTIX_SCENE_MAX_ENTRIES 256  /* part pairs over all groups (sum of n_g)                  */
typedef struct CotixScene CotixScene;
typedef struct CotixSceneParams {
  float dt;                /* 1e-2 */
  float init_dir[2];       /* random_direction(PRNGKey(1)), cotix/_collisions.py:287-298 */
  int32_t gjk_max_iters;   /* 32 */
  float coin_p;            /* 0.5, cotix/_colliders.py:222-224 */
  float baumgarte;         /* 0.3 */
  float baumgarte_dt;      /* 0.01 */
  int32_t advance_key;     /* 1: keys <- split(key)[0] after the step */
} CotixSceneParams;
void cotix_scene_default_params(CotixSceneParams* p);
/*
 * Static analysis of a scene (host) and its upload to `device` (device < 0: host analysis only,
 * for cotix_scene_info).  body_props [nb][4] = mass, inertia, elasticity, friction_coefficient
 * (cotix/_bodies.py:189-206; inf mass = static); part_body[nparts] = owning body, parts of a
 * body in the order of its UniversalShape; part_recs [nparts][16] = LOCAL shape records.  Fails
 * with COTIX_EINVAL where the reference raises: a pair of part types that is not a key of
 * _contact_funcs in (higher body index, lower body index) order (cotix/_colliders.py:99-111).
 */
int cotix_scene_create(CotixScene** out, int device, int32_t nb, const float* body_props,
                       int32_t nparts, const int32_t* part_body, const float* part_recs);
int cotix_scene_destroy(CotixScene* scene);
/* groups (type pairs in first-appearance order), distinct evaluations per step, and the
 * sum of n_g^2 evaluations the reference's vmap-squared makes (cotix/_colliders.py:149-173) */
int cotix_scene_info(const CotixScene* scene, int32_t* n_groups, int32_t* n_pairs,
                     int64_t* n_reference_evals, int32_t* group_sizes /*[<= 9] or NULL*/);
/*
 * state [nb * 6][ld] (DEVICE, in place): px, py, vx, vy, angle, angular_velocity of body b at
 * rows 6b .. 6b+5; keys [2][ld] uint32 (DEVICE); optional outputs: cell [nb * nb][ld] (live
 * cells of all_contacts after the coin flips), chosen [nb][ld] (contact_index per body),
 * status [ld] (COTIX_ST_* bits).  Enqueues one kernel on `stream`; never syncs or allocates.
 */
int cotix_scene_step(void* stream, const CotixScene* scene, int64_t B, float* state, int64_t ld,
                     uint32_t* keys, const CotixSceneParams* p, uint8_t* cell, int32_t* chosen,
                     uint32_t* status);

/* ------------------------------------------------------------------------------------ */
/* Small primitives exposed for parity tests of the reference's leaf functions.           */
/* ------------------------------------------------------------------------------------ */
/* ExplicitEulerPhysics._single_body_step, cotix/_physics_solvers.py:17-24; arrays [n]    */
int cotix_euler_step(void* stream, int64_t n, float* px, float* py, const float* vx,
                     const float* vy, float* ang, const float* angvel, float dt);
/* resolve_collision, cotix/_collision_resolution.py:52-151; bodies as [n][10] floats in   */
/* DynamicBody leaf order (mass, inertia, px, py, vx, vy, angle, angvel, elasticity,       */
/* friction), contact as pen[n][2], cp[n][2]; updated in place                             */
int cotix_resolve_collision(void* stream, int64_t n, float* body1, float* body2,
                            const float* pen, const float* cp, float baumgarte,
                            float baumgarte_dt);
/* get_support for a batch of (record, direction) pairs -> out[n][2];                      */
/* cotix/_convex_shapes.py:22-26,62-66,149-155                                             */
int cotix_support(void* stream, int64_t n, const float* recs, const float* dirs,
                  float* out);
/* jax.random primitives on the device, for known-answer tests of the RNG restatement:    */
/* out[i] = threefry2x32(key, (x0[i], x1[i])) ; out is [n][2]                              */
int cotix_threefry2x32(void* stream, int64_t n, uint32_t k0, uint32_t k1,
                       const uint32_t* x0, const uint32_t* x1, uint32_t* out);
/* out[i] = philox4x32_10(key, ctr[i]) ; ctr, out are [n][4]                               */
int cotix_philox4x32(void* stream, int64_t n, uint32_t k0, uint32_t k1,
                     const uint32_t* ctr, uint32_t* out);

/* ------------------------------------------------------------------------------------ */
/* Device memory, streams and events for hosts that have no CUDA binding of their own      */
/* (ctypes, cgo, JNI ...): thin wrappers over the CUDA runtime so that the Python mirror    */
/* runs without torch.  `stream` NULL = the default stream.  cotix_memcpy infers the        */
/* direction from the pointers (cudaMemcpyDefault) and is asynchronous on `stream` for      */
/* page-locked host memory.                                                                 */
/* ------------------------------------------------------------------------------------ */
int cotix_set_device(int device);
int cotix_device_count(int* n);
int cotix_device_malloc(int64_t bytes, void** out);
int cotix_device_free(void* p);
int cotix_host_malloc(int64_t bytes, void** out); /* page-locked */
int cotix_host_free(void* p);
int cotix_memcpy(void* stream, void* dst, const void* src, int64_t bytes);
int cotix_memset(void* stream, void* dst, int value, int64_t bytes);
int cotix_stream_create(void** out);
int cotix_stream_destroy(void* stream);
int cotix_stream_synchronize(void* stream);
int cotix_event_create(void** out);
int cotix_event_destroy(void* event);
int cotix_event_record(void* event, void* stream);
int cotix_event_elapsed_ms(void* start, void* stop, float* ms); /* waits for `stop` */

/* ------------------------------------------------------------------------------------ */
const char* cotix_last_error(void);
int cotix_abi_version(void);
/* 0: the exact build (-fmad=false, IEEE division / sqrt: bit-identical to the CPU oracle);          */
/* 1: libcotix_b200_fast.so, the same sources built -fmad=true -prec-div=false -prec-sqrt=false:   */
/*    within 1e-5 relative per step of the exact build, not bit-exact         */
int cotix_math_flavour(void);
/* number of kernels this library has launched in this process (all streams) */
int64_t cotix_launch_count(void);
/* name, SM count and compute capability of `device`; returns 0 or COTIX_E*           */
int cotix_device_info(int device, char* name, int name_len, int* sm_count, int* cc_major,
                      int* cc_minor);

#ifdef __cplusplus
}
#endif
#endif /* COTIX_B200_H */

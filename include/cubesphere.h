/*
 * cubesphere.h -- C ABI of the B200-native cubed-sphere shallow-water hot path.
 *
 * One explicit time step over the six gnomonic-equiangular panels: halo fill,
 * finite-volume flux/tendency stencil, fused Runge-Kutta stage update, and the
 * mass / energy reductions, as hand-written sm_100a CUDA kernels.
 *
 * Boundary rules
 *   - plain C: pointers, sizes, PODs; no C++ or torch types cross this header;
 *   - every device pointer is caller-owned memory on the context's device;
 *   - the compute calls (cs_halo_fill_*, cs_halo_pack / unpack, cs_rk_stage[_part], cs_step,
 *     cs_diagnostics[_noalloc]) only ENQUEUE work on the given stream (a cudaStream_t passed as
 *     void*; NULL = the legacy default stream); they neither synchronise nor allocate.  One
 *     exception: the first cs_step for a new (h, u, v, dt, diag_out) combination captures and
 *     instantiates a CUDA graph (creates a stream, allocates graph memory; CS_STEP_GRAPH=0 opts out);
 *   - the set-up calls synchronise the device and/or allocate: cs_ctx_create, cs_ctx_destroy,
 *     cs_halo_set_rotation, cs_halo_set_interp, cs_halo_interp_sharded, cs_halo_plan, cs_halo_peer_alloc, cs_halo_peer_connect, and cs_step_host[_batch] (which own their copies);
 *   - every call runs on the context's device and restores the calling thread's current CUDA
 *     device before it returns;
 *   - return value 0 = success, negative = CS_E_* below; the text of the last
 *     error of a context (or of the calling thread when ctx is NULL) is
 *     cs_last_error();
 *   - there is no CPU fallback: without a CUDA device every compute entry point
 *     fails with CS_E_CUDA.
 *
 * State layout (reference JAX-DevLab-Examples.py:141,174; deck slide 11):
 *   each field is one dense array  [tile_count][n + 2g][n + 2g], x fastest,
 *   n = N / tiles_per_edge, interior cells at [g, g + n) on both axes, tile id =
 *   face * p^2 + ty * p + tx.  A context owns the contiguous range of POSITIONS
 *   [tile_begin, tile_begin + tile_count) of axis 0 -- the block partition that
 *   the reference gets from NamedSharding(mesh, P('tiles'))
 *   (JAX-DevLab-Examples.py:77-79); position k holds tile k unless cs_config.tile_order
 *   says otherwise.  h is the fluid depth (m); u, v are the
 *   panel-local contravariant velocity components (rad/s).
 *
 * "Replaces" below cites the reference interface each entry point stands in for
 * (file /root/reference/JAX-DevLab-Examples.py, abbreviated PY; deck = the PDF).
 */
#ifndef CUBESPHERE_H_
#define CUBESPHERE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CS_ABI_VERSION 1

#if defined(__GNUC__)
#define CS_API __attribute__((visibility("default")))
#else
#define CS_API
#endif

/* error codes */
#define CS_OK 0
#define CS_E_INVALID (-1)        /* bad argument / configuration (Python: ValueError)        */
#define CS_E_UNSUPPORTED (-2)    /* valid but not implemented (Python: NotImplementedError)  */
#define CS_E_CUDA (-3)           /* a CUDA runtime call failed; see cs_last_error            */
#define CS_E_NOMEM (-4)          /* device or host allocation failed                         */

/* enums */
#define CS_F32 0
#define CS_F64 1
#define CS_RK_EULER 1
#define CS_RK_SSPRK2 2
#define CS_RK_SSPRK3 3
#define CS_LIM_NONE 0
#define CS_LIM_MINMOD 1
#define CS_LIM_MC 2
#define CS_SUMS_OF_OUTPUT 0      /* cs_rk_stage_part_sums: sums over qout                    */
#define CS_SUMS_OF_INPUT 1       /*                        sums over qs (the stage input)    */
#define CS_KERNEL_AUTO 0         /* fastest validated stage kernel                           */
#define CS_KERNEL_NAIVE 1        /* one thread per cell, global loads only (cross-check)     */
#define CS_KERNEL_MARCH 2        /* register-marching kernel, direct global loads            */
#define CS_KERNEL_TMA 3          /* register-marching kernel fed by a TMA shared-memory ring */

typedef struct cs_ctx cs_ctx;

/* Versioned POD.  Zero-initialise, fill, pass to cs_ctx_create. */
typedef struct cs_config {
    int32_t abi_version;      /* CS_ABI_VERSION                                              */
    int32_t N;                /* cells per face edge (C48 -> 48)                             */
    int32_t tiles_per_edge;   /* p; 6 p^2 tiles (PY:40).  reference supports 1 only (PY:31)  */
    int32_t halo;             /* g; the step needs g >= 2, plain halo fills accept g >= 1    */
    int32_t dtype;            /* CS_F32 | CS_F64                                             */
    int32_t rk;               /* CS_RK_*                                                     */
    int32_t limiter;          /* CS_LIM_*                                                    */
    int32_t advect_only;      /* 1: velocity frozen (Williamson case 1)                      */
    int32_t device;           /* CUDA device ordinal                                         */
    int32_t tile_begin;       /* first tile owned by this context                            */
    int32_t tile_count;       /* number of tiles owned; 0 = all 6 p^2                        */
    int32_t kernel;           /* CS_KERNEL_*                                                 */
    int32_t fused_halo;       /* 1: stage kernel also stores boundary cells into the neighbours' edge ghosts  */
                              /*    (no halo launch; corner blocks not refreshed).  Off by default: measured  */
                              /*    slower than the separate 3-microsecond halo launch   */
    int32_t sums_of_entry_state; /* 0: cs_step's diag_out = mass / energy of the NEW state (summed inside the last stage
                                  *    launch); 1: of the state the step STARTS from (summed inside the first stage launch,
                                  *    where it costs a few instructions per cell and no registers: state n's sums arrive
                                  *    with step n + 1, and a final cs_diagnostics gives the last state's) */
    const int32_t* tile_order;/* optional host table of 6 p^2 entries: position along axis 0 -> tile id (face p^2 + ty p + tx), */
                              /* a permutation.  tile_begin / tile_count (and the peers of cs_halo_plan) then count POSITIONS, */
                              /* so a block partition of axis 0 can hand each device a spatially compact set of tiles          */
                              /* (cubesphere.partition.locality_order).  NULL -> identity.  (Occupies the former reserved[2].)  */
    double radius;            /* a      (0 -> 6.37122e6 m)                                   */
    double omega;             /* Omega  (0 -> 7.292e-5 1/s; set advect_only for no rotation) */
    double gravity;           /* g      (0 -> 9.80616 m/s^2)                                 */
    double axis[3];           /* unit rotation axis (all 0 -> (0,0,1))                       */
    const double* tangents;   /* optional host table of N centre tangents tan(alpha_i);      */
                              /* NULL -> computed internally                                 */
} cs_config;

/* Library version of this ABI.  Replaces: nothing (the reference has no FFI). */
CS_API int cs_abi_version(void);

/* Text of the last error of ctx, or of the calling thread when ctx == NULL. */
CS_API const char* cs_last_error(const cs_ctx* ctx);

/*
 * Validate the configuration, select the device, build the seam gather tables
 * and metric tables, allocate the Runge-Kutta scratch state.
 * Replaces: setup_sharding's validation of tiles_per_edge / divisibility
 * (PY:31-57), create_communication_schedule (PY:105-139) and the factory
 * make_halo_exchange(schedule, N) (PY:199-246), whose frozen static arguments
 * (faces, edges, ops, N) become the context's tables.
 */
CS_API int cs_ctx_create(const cs_config* cfg, cs_ctx** out);
CS_API int cs_ctx_destroy(cs_ctx* ctx);

/*
 * Fill every ghost cell (4 edge strips of width g and 4 g x g corner blocks per
 * tile) of one scalar field in place.  Bit-exact copy of neighbour interiors
 * with the seam's index reversal.
 * Replaces: cubesphere_halo_exchange(field_ghosts) (PY:238-242) = 12 x
 * exchange_edge_pair (PY:166-197) incl. extract_boundary_data / apply_operations
 * / set_ghost_data (PY:143-163,184-195); ghost corners, which the reference never
 * writes, get the rule of DESIGN.md section 2.4.
 * On a sharded context (tile_count < 6 p^2) only the ghosts whose source tile is owned by the same
 * context are filled; the others come through cs_halo_pack / cs_halo_unpack.
 */
CS_API int cs_halo_fill_scalar(cs_ctx* ctx, void* field, void* stream);

/*
 * Same for a panel-local vector field: copy + index reversal + the seam's signed
 * permutation of the two components (bit-exact).
 * Replaces: the reference's "Cartesian Velocity Exchange" (deck slide 18), which
 * side-steps the rotation by exchanging three Cartesian scalars.
 */
CS_API int cs_halo_fill_vector(cs_ctx* ctx, void* u, void* v, void* stream);

/* h, u, v in one launch.  Replaces: three calls of the above per RK stage. */
CS_API int cs_halo_fill_state(cs_ctx* ctx, void* h, void* u, void* v, void* stream);

/*
 * Optional position-dependent 2x2 rotation of vector ghosts: rot is a HOST array
 * [ghost cells][4] (row-major 2x2, fp64) in the order of cs_halo_table_info;
 * ghost (u, v) = R (u, v)_neighbour, evaluated as two products and one sum per
 * component without FMA contraction, so a CPU evaluation in the same order is
 * bit-identical.  NULL restores the signed permutation.
 */
CS_API int cs_halo_set_rotation(cs_ctx* ctx, const double* rot, size_t n_ghost);

/*
 * Optional second-order ghosts: instead of copying the neighbour's cell, every ghost cell becomes the value at the
 * point of OUR extended grid, ((w0 s0 + w1 s1) + w2 s2) + w3 s3 over four interior cells of the neighbour (1-D Lagrange
 * interpolation along its row at the same depth), vectors then through the 2 x 2 `rot` at that point -- each product
 * and sum rounded on its own, so a CPU evaluation in the same order is bit-identical.  Host tables, one row per ghost
 * cell, any order: dst [n], src [n][4] (element offsets into a field), w [n][4], rot [n][4] row-major, all fp64;
 * tangents (may be NULL) = N + 2 halo tangents replacing the context's periodic table, so that ghost cells carry the
 * metric of the extended chart.  cubesphere.geometry.seam_interp_tables builds all five.  dst == NULL restores the
 * plain copy and the periodic metric.  Whole-sphere contexts only.  Trade-off: second order at the
 * seams instead of first, but seam fluxes are no longer single valued, so mass drifts at truncation level
 * (1e-7 per day at C96) instead of round-off.  Synchronises the device; allocates.
 * Replaces: nothing in the reference (its halo is the plain copy, PY:183-195); the deck's "solid 2nd-order method"
 * (slide 4) is what this option delivers across seams.
 */
CS_API int cs_halo_set_interp(cs_ctx* ctx, size_t n_ghost, const int64_t* dst, const int64_t* src, const double* w,
                       const double* rot, const double* tangents);

/*
 * The same second-order ghosts on a SHARD (tile_count < 6 p^2), with the one-sided exchange.  The four sources of a ghost
 * cell may live on this context or on peers, so raw cells travel and the receiver interpolates: send_counts[r] /
 * send_src (concatenated, element offsets into the local shard) = the cells peer r reads, in the order it expects them;
 * recv_counts[r] = the cells of peer r this context's rows read; rows as in cs_halo_set_interp, except that src addresses
 * one virtual space -- offsets below the shard's element count are local cells, shard_elems + k is cell k of the receive
 * buffer (peer blocks in rank order).  cubesphere.partition.InterpExchangePlan derives all of it from the global tables,
 * consistently on every rank.  Call it in place of cs_halo_plan, then cs_halo_peer_alloc / cs_halo_peer_connect as usual;
 * afterwards cs_halo_push + cs_halo_pull produce EVERY ghost cell of the shard (cs_halo_fill_* becomes a no-op).
 * Replaces: nothing in the reference (see cs_halo_set_interp).
 */
CS_API int cs_halo_interp_sharded(cs_ctx* ctx, int n_peers, const int64_t* send_counts, const int64_t* send_src,
                           const int64_t* recv_counts, size_t n_rows, const int64_t* dst, const int64_t* src, const double* w,
                           const double* rot, const double* tangents);

/*
 * Ghost-cell table of the context, for callers that want to build rotation
 * tables: writes up to cap entries of (dst_offset, src_offset, code) and returns
 * the ghost-cell count in *n_ghost.  Offsets are element offsets into a field of
 * the local shard; src_offset is -1 for cells whose source lives on another
 * context.  Any output pointer may be NULL.
 */
CS_API int cs_halo_table_info(const cs_ctx* ctx, size_t cap, int64_t* dst_offset, int64_t* src_offset,
                       int32_t* code, size_t* n_ghost);

/*
 * Static bottom topography hs (same layout and dtype as the state); interior
 * cells are copied into the context and its ghosts filled.  NULL clears it.
 */
CS_API int cs_set_topography(cs_ctx* ctx, const void* hs, void* stream);

/*
 * One fused Runge-Kutta stage on all owned tiles:
 *     qout = a * q0 + b * (qs + dt * L(qs))      on interior cells,
 * L = PLR-upwind continuity + vector-invariant momentum with Coriolis and
 * pressure-gradient sources.  qs must be halo-filled; qout's ghosts are filled
 * on return (fused push or a halo launch).  qout may alias q0, never qs.
 * When a == 0, q0 is not read.  diag_out (device double[2], may be NULL) receives
 * mass and total energy of qout.
 * Replaces: the "Numerics (FV)" box driven by jax.shard_map (deck slides 6-7);
 * not shipped in the reference.
 */
CS_API int cs_rk_stage(cs_ctx* ctx, const void* const q0[3], const void* const qs[3], void* const qout[3],
                double a, double b, double dt, double* diag_out, void* stream);

/*
 * One explicit time step, in place: h, u, v must be halo-filled on entry and are
 * halo-filled on return.  Replaces: step(state, dt) of BASELINE.json's
 * north_star (the solver driver behind deck slide 7; not shipped).
 * diag_out (device double[2], may be NULL): mass and total energy of the new state, or -- on a context created
 * with sums_of_entry_state = 1 -- of the state on entry.  Either way they come out of a stage launch, not out of
 * another pass over the state.
 */
CS_API int cs_step(cs_ctx* ctx, void* h, void* u, void* v, double dt, double* diag_out, void* stream);

/*
 * Demo workload (deck slides 12, 17: heat diffusing from a checkerboard source on the top panel, "proof that
 * sharding works"): one explicit step of scalar diffusion, Tout = T + kappa_dt * L(T) + dt_source * source on the
 * interior cells, L = the 5-point flux-form Laplacian with the panel metric (no cross-derivative terms: those need
 * the corner ghosts no panel owns), then Tout's ghosts are filled.  T must be ghost-filled; source may be NULL; same
 * layout and dtype as the state.  Total heat sum sqrt(G) T is conserved to round-off (single-valued seam fluxes).
 * Stable for kappa_dt <= delta^2 / (4 max g^ii).  Replaces: the thermal-diffusion demo solver (not shipped upstream).
 */
CS_API int cs_scalar_diffuse(cs_ctx* ctx, const void* T, void* Tout, const void* source, double kappa_dt, double dt_source,
                      void* stream);

/*
 * Mass  M = sum sqrt(G) h dalpha dbeta  and total energy
 * E = sum sqrt(G) (h K + g h^2 / 2 + g h hs) dalpha dbeta  over the owned interior
 * cells, accumulated in fp64 in a fixed order (deterministic); out = device
 * double[2].  Replaces: conservation diagnostics named by north_star; not shipped.
 */
CS_API int cs_diagnostics(cs_ctx* ctx, const void* h, const void* u, const void* v, double* out, void* stream);

/*
 * Host-buffer convenience entry (what a ctypes / cgo style caller with NumPy
 * arrays uses): copies h, u, v (host, full padded layout) to the device, fills
 * halos, runs nsteps steps, writes mass/energy of the final state to diag_host[2]
 * (may be NULL) and, when copy_back != 0, copies the state back.  Synchronises.
 */
CS_API int cs_step_host(cs_ctx* ctx, void* h_host, void* u_host, void* v_host, double dt, int nsteps,
                 double* diag_host, int copy_back);

/*
 * The same for `members` independent host-resident states (an ensemble, or a
 * batch of requests): member i is h_host[i], u_host[i], v_host[i]; its mass and
 * energy go to diag_host[2 i .. 2 i + 1] (diag_host may be NULL).  The members
 * stream through two device slots on three internal streams, so the upload of
 * member i + 1 and the download of member i - 1 overlap the steps of member i;
 * use page-locked host memory or the copies serialise.  Results are those of
 * `members` cs_step_host calls.  Synchronises.  No counterpart in the reference
 * (its states live on the devices, PY:174); this is the host-array form of
 * step(state, dt) for more states than fit in, or are wanted on, the GPU.
 */
CS_API int cs_step_host_batch(cs_ctx* ctx, int members, void* const* h_host, void* const* u_host, void* const* v_host,
                       double dt, int nsteps, double* diag_host, int copy_back);

/*
 * Sharded halo exchange (tile_count < 6 p^2).  The context splits its ghost
 * cells into local ones (filled by cs_halo_fill_* / the stage kernel) and remote
 * ones.  For each peer context r (in order of tile_begin) cs_halo_plan reports
 * how many cells this context sends to and receives from r; cs_halo_pack gathers
 * the boundary cells of h, u, v that peers need into sendbuf (3 values per cell,
 * peers in rank order), the caller moves the bytes (NCCL send/recv, all_to_all or
 * peer copies), and cs_halo_unpack scatters recvbuf into the ghosts applying the
 * seam permutation.  nfields = 1 packs a single scalar field (h only).
 * Replaces: the cross-device copies GSPMD inserts for .at[face].set(...) on an
 * array sharded with P('tiles') (PY:79,192-195).
 */
CS_API int cs_halo_plan(cs_ctx* ctx, int n_peers, const int32_t* peer_tile_begin,
                 const int32_t* peer_tile_count, int64_t* send_counts, int64_t* recv_counts);
CS_API int cs_halo_pack(cs_ctx* ctx, const void* h, const void* u, const void* v, int nfields,
                 void* sendbuf, void* stream);
CS_API int cs_halo_unpack(cs_ctx* ctx, const void* recvbuf, void* h, void* u, void* v, int nfields,
                   void* stream);

/*
 * The same exchange with one-sided peer stores over NVLink (GPUs of one box, one process per GPU) instead of
 * pack -> all-to-all -> unpack.  Set-up, once, after cs_halo_plan:
 *   cs_halo_peer_alloc   allocates this context's double-buffered receive buffer and flag words and writes their
 *                        CUDA IPC handles (CS_PEER_HANDLE_BYTES bytes) to handles_out;
 *   (the caller gathers every rank's handle block and receive counts, e.g. with torch.distributed.all_gather)
 *   cs_halo_peer_connect maps the peers' buffers: all_handles = [n_peers][CS_PEER_HANDLE_BYTES] in rank order,
 *                        remote_offsets[r] = the offset, in cells, of this rank's block inside rank r's receive
 *                        buffer (the sum of r's receive counts from the ranks before this one).
 * Per ghost fill, on every rank: cs_halo_push (gathers the boundary cells each peer needs and stores them straight
 * into that peer's receive buffer, then raises the peer's flag for this rank) and cs_halo_pull (waits for the flags
 * of this fill, scatters the buffer into the remote ghosts with the seam permutation).  Both only enqueue; fills
 * are counted on the device, so the pair can be captured into a CUDA graph and replayed.  Every rank must issue
 * the same sequence of fills; destroy the contexts only after all ranks have finished (the peers hold mappings).
 * Replaces: the cross-device copies GSPMD inserts for .at[face].set(...) (PY:79,192-195); north_star names
 * CUDA-IPC peer stores first, NCCL send/recv second.
 */
#define CS_PEER_HANDLE_BYTES 192
CS_API int cs_halo_peer_alloc(cs_ctx* ctx, void* handles_out, size_t handles_bytes);
CS_API int cs_halo_peer_connect(cs_ctx* ctx, int n_peers, int my_rank, const void* all_handles, const int64_t* remote_offsets);
/* The same connection for contexts of ONE process (several devices with peer access enabled, or several contexts on one
 * device): no IPC handles -- cs_halo_peer_pointers gives the three device pointers of a context's buffers (out3[3]) and
 * cs_halo_peer_connect_pointers takes everybody's, [n_peers][3] in rank order.  Same rules otherwise. */
CS_API int cs_halo_peer_pointers(cs_ctx* ctx, void** out3);
CS_API int cs_halo_peer_connect_pointers(cs_ctx* ctx, int n_peers, int my_rank, void* const* all_pointers,
                                  const int64_t* remote_offsets);
CS_API int cs_halo_push(cs_ctx* ctx, const void* h, const void* u, const void* v, int nfields, void* stream);
CS_API int cs_halo_pull(cs_ctx* ctx, void* h, void* u, void* v, int nfields, void* stream);
/* How many waits of a pull gave up before a peer's flag arrived (option "peer_timeout_ms", default 5000; the ghosts of
 * that fill are then stale).  Non-zero means the ranks' fills fell out of step or a peer died; the failure is sticky:
 * once a wait has given up, no later pull of this context waits, so a broken job drains instead of paying the time-out
 * on every fill.  Callers check this where they synchronise anyway (the Python mirror raises).  Synchronises. */
CS_API int cs_halo_peer_timeouts(cs_ctx* ctx, unsigned int* count_host);

/*
 * Split stage for overlap on shards: part = 0 computes cells that need no remote
 * ghost, part = 1 the rim that does, part = 2 everything.  Same contract as
 * cs_rk_stage otherwise (no halo fill of qout is done for parts 0 / 1).
 */
CS_API int cs_rk_stage_part(cs_ctx* ctx, const void* const q0[3], const void* const qs[3], void* const qout[3],
                     double a, double b, double dt, int part, void* stream);

/*
 * cs_rk_stage_part that also leaves, in sums_out (device double[2], may be NULL), the mass and total energy summed
 * over the cells THIS part computes (part 0 + part 1 = the shard; add the two, then add over shards): of the new
 * values (which = CS_SUMS_OF_OUTPUT) or of the stage input qs at those cells (CS_SUMS_OF_INPUT, much cheaper).  The
 * sums come out of the stage launch itself: nothing is read twice.  Available for the marching kernel (parts 0, 2) and
 * the rim kernel (part 1); other variants return CS_E_UNSUPPORTED after enqueuing the stage.  Parts 0 and 1 may
 * run at the same time on two streams (they use separate scratch); give them separate sums_out.
 * Replaces: conservation diagnostics named by north_star, for a state sharded with P('tiles') (PY:77-79).
 */
CS_API int cs_rk_stage_part_sums(cs_ctx* ctx, const void* const q0[3], const void* const qs[3], void* const qout[3],
                          double a, double b, double dt, int part, int which, double* sums_out, void* stream);

/*
 * Mass and total energy like cs_diagnostics, for callers that cannot allocate device memory and
 * therefore cannot create a cs_ctx: plugin handlers inside a framework that has reserved the GPU
 * for its own allocator (under JAX/XLA preallocation a cudaMalloc from a plugin can fail), and
 * monitoring hooks on latency-critical threads (cudaMalloc / cudaFree lock the device and may wait
 * for running work).  No counterpart in the reference; no context; calls no cudaMalloc, cudaFree or
 * any allocating routine.  The caller passes the inputs (h, u, v, optional hs: same layout as the
 * state, dtype = CS_F32 | CS_F64, for the tile range [tile_begin, tile_begin + tile_count),
 * tile_count 0 = all) and the output (device double[2]).  radius / gravity 0 select the defaults.
 * Scratch is a fixed static array inside the library (about 128 KB on every GPU that loads it, kept
 * until exit): at most 33,554,432 interior cells per call (C1536 whole sphere: 14,155,776), larger
 * shards return CS_E_INVALID.  The metric is that of the default equiangular grid: a custom tangent
 * table (cs_config.tangents) is a context feature and is NOT honoured here.  The calling thread's
 * current device is restored on return.  The first call on a device resolves the array's address; make that
 * call before any stream capture on that device.  All calls on a device share the array (it keeps
 * the last call's partial sums; the ticket counter is cleared on every call), so calls whose
 * kernels may overlap on the GPU race: serialise them on one stream.
 */
CS_API int cs_diagnostics_noalloc(int dtype, int N, int tiles_per_edge, int halo, int tile_begin, int tile_count,
                                  double radius, double gravity, const void* h, const void* u, const void* v,
                                  const void* hs, double* out, int device, void* stream);

/*
 * Tuning knobs of one context, by name (the CS_* environment variables of DESIGN.md section 4 set the same knobs
 * for every context at creation): "march_prefetch", "march_rows", "march_vec", "march_unit", "reserve_sms",
 * "tma_ring", "fused_halo", "step_graph", "fused_diag", "step_kernel", "peer_timeout_ms".  Drops the context's captured step graphs.
 * Replaces: nothing (tuning).
 */
CS_API int cs_ctx_set_option(cs_ctx* ctx, const char* name, int value);

/*
 * Which stage kernel ran: *last_variant = CS_KERNEL_* of the most recent cs_rk_stage / cs_step launch
 * (CS_KERNEL_AUTO before the first one), *naive_fallbacks = how many launches a CS_KERNEL_AUTO context had to
 * send to the one-thread-per-cell kernel (about 4x slower) because the fields were not 16-byte aligned, the row
 * length was odd or halo < 2.  Either pointer may be NULL.  Replaces: nothing (bookkeeping).
 */
CS_API int cs_stage_kernel_info(const cs_ctx* ctx, int32_t* last_variant, int64_t* naive_fallbacks);

/*
 * Tuning aid.  With CS_STEP_DEBUG=16 in the environment when the context is created, the one-launch-per-step kernel
 * records nanosecond time stamps per thread block (start; end of each stage's march; release from each grid barrier;
 * end of the final ghost fill): out_host receives [max_blocks][8] values of the last launch.  Synchronises the device.
 * Replaces: nothing (profiling).
 */
CS_API int cs_step_kernel_trace(cs_ctx* ctx, int max_blocks, unsigned long long* out_host);

/* Number of kernels this library has launched on ctx since creation (bench bookkeeping). */
CS_API int64_t cs_launch_count(const cs_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* CUBESPHERE_H_ */

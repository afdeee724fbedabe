/* sphexa_b200.h — C ABI of the B200-native (sm_100a) SPH-EXA hot path.
 *
 * The reference (lks1248/SPH-EXA) has no C ABI: its GPU boundary is a set of
 * `extern` C++ function templates explicitly instantiated in .cu files and
 * archived as cstone_gpu / sph_gpu.  Every entry point below names the
 * reference interface it replaces (file:line relative to the reference root).
 * A maintainer binds these from the reference's accelerator switch points
 * (`if constexpr (HaveGpu<Acc>{})`), see INTEGRATION.md.
 *
 * Conventions
 *  - all pointers are DEVICE pointers unless the name ends in _h / says "host"
 *  - every call returns 0 on success, non-zero on error; sphb200_last_error()
 *    gives the message (the C++ shim rethrows std::runtime_error like the reference)
 *  - all work is issued on the context's stream; calls that return scalars to the
 *    host synchronise that stream, all others are asynchronous
 *  - KeyType = uint64_t (63-bit Hilbert keys), LocalIndex = uint32_t,
 *    TreeNodeIndex = int32_t  (domain/include/cstone/tree/definitions.h:42)
 *  - "real" fields: void* interpreted by `dtype`: SPHB200_F64 -> double (north-star
 *    parity mode, T=Tm=double like sph/test/ve.cpp:53), SPHB200_MIXED -> float
 *    (reference default HydroType, sph/include/sph/types.hpp:39-46).
 *    Coordinates x,y,z, temp and du are always double (CoordinateType).
 */
#ifndef SPHEXA_B200_H
#define SPHEXA_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SPHB200_F64 0
#define SPHB200_MIXED 1

#define SPHB200_OPEN 0
#define SPHB200_PERIODIC 1
#define SPHB200_FIXED 2

/* mirrors cstone::Box<double> (domain/include/cstone/sfc/box.hpp:111-191) */
typedef struct sphb200_box
{
    double lim[6]; /* xmin xmax ymin ymax zmin zmax */
    double len[3]; /* xmax-xmin ...                 */
    double inv[3]; /* 1/(xmax-xmin) ...             */
    int bc[3];     /* SPHB200_OPEN / PERIODIC / FIXED */
} sphb200_box;

/* mirrors cstone::OctreeNsView<double,uint64_t> (domain/include/cstone/tree/octree.hpp:295-316) */
typedef struct sphb200_tree_view
{
    int numLeafNodes;
    int numNodes;
    const uint64_t* prefixes;
    const int32_t* childOffsets;
    const int32_t* internalToLeaf;
    const int32_t* levelRange;
    const uint64_t* leaves;
    const uint32_t* layout;
    const double* centers; /* [numNodes][3] */
    const double* sizes;   /* [numNodes][3] */
} sphb200_tree_view;

/* particle field table, mirrors the device half of sph::ParticlesData
 * (sph/include/sph/particles_data.hpp:203-232, particles_data_gpu.cuh) */
typedef struct sphb200_fields
{
    /* conserved */
    double *x, *y, *z;
    void *h, *m, *vx, *vy, *vz, *x_m1, *y_m1, *z_m1;
    double* temp;
    void *du_m1, *alpha;
    /* dependent */
    void *xm, *kx, *gradh, *prho, *c;
    void *c11, *c12, *c13, *c22, *c23, *c33;
    void *divv, *curlv;
    void *ax, *ay, *az;
    double* du;
    uint32_t* nc;                               /* neighbor count + 1 (self), indexed by particle index */
    void *dV11, *dV12, *dV13, *dV22, *dV23, *dV33; /* only with avClean, else NULL */
    size_t n; /* length of every array (owned + halo particles); neighbor indices are < n */
} sphb200_fields;

/* physics constants, mirrors the scalar members of sph::ParticlesData
 * (sph/include/sph/particles_data.hpp:87-138) */
typedef struct sphb200_hydro_const
{
    double K;        /* kernel normalisation (kernel_3D_k) */
    unsigned ng0, ngmax;
    double Kcour, Krho;
    double gamma, muiConst;
    double alphamin, alphamax, decay_constant;
    double Atmin, Atmax, ramp;
    double minDt, minDt_m1;
    const void* wh;  /* device, 20000 entries of `real` */
    const void* whd; /* device, 20000 entries of `real` */
    int avClean;
} sphb200_hydro_const;

typedef struct sphb200_ctx sphb200_ctx;

/* ---- context ---- */
/* stream: a cudaStream_t (e.g. torch.cuda.current_stream().cuda_stream; NULL = the legacy default stream) or
 * SPHB200_PRIVATE_STREAM to let the context create its own non-blocking stream */
#define SPHB200_PRIVATE_STREAM ((void*)(intptr_t)-1)
int sphb200_create(int device, void* stream, sphb200_ctx** out);
void sphb200_destroy(sphb200_ctx* ctx);
const char* sphb200_last_error(void);
int sphb200_sync(sphb200_ctx* ctx);
/* number of kernels this library launched since creation (bench.py gpu_launches) */
uint64_t sphb200_launch_count(const sphb200_ctx* ctx);
/* host helper: fill len/inv like the cstone::Box constructor (box.hpp:128-141) */
void sphb200_make_box(const double lim_h[6], const int bc_h[3], sphb200_box* out);

/* The transition table behind sphb200_sfc_keys (host copy, for inspection and tests; no device needed): 24 states x 512
 * inputs, entry = 9 key bits | next state << 9 for the three levels whose coordinate bits are x3 << 6 | y3 << 3 | z3,
 * stored at the bank-swizzled position idx ^ ((idx >> 5) & 0x3E) of idx = state * 512 + input.  Returns the number of
 * states, -1 on error. */
int sphb200_hilbert_table(uint16_t* table_h, size_t capacity);
/* ---- a2: computeSfcKeysGpu  (domain/include/cstone/sfc/sfc_gpu.h:39, sfc_gpu.cu:38-55) ---- */
int sphb200_sfc_keys(sphb200_ctx*, const double* x, const double* y, const double* z, size_t n, const sphb200_box* box_h,
                     uint64_t* keys);
/* ---- §8f row 4: initial conditions from a glass (template) block, main/src/init/grid.hpp ----
 * assembleCuboid (grid.hpp:238-264): the global box is tiled by multiplicity[0] x [1] x [2] copies of a template block
 * whose coordinates lie in [0,1)^3 (device arrays of blockSize); the call writes the copies that fall into the octree
 * cells spanning the SFC range [keyStart, keyEnd) (a rank's share: equal slices of [0, 2^63)), in the reference's
 * order (cell, tile, position in the block), and returns their number in *count_h.  SPHB200_ERR_CAPACITY with the
 * needed count in *count_h when x, y, z (capacity elements each) are too small; nothing is written then. */
int sphb200_assemble_cuboid(sphb200_ctx*, uint64_t keyStart, uint64_t keyEnd, const sphb200_box* globalBox_h,
                            const int multiplicity_h[3], const double* xBlock, const double* yBlock,
                            const double* zBlock, size_t blockSize, double* x, double* y, double* z, size_t capacity,
                            uint64_t* count_h);
/* the octree cells sphb200_assemble_cuboid works through: cstone::spanSfcRange (sfc/common.hpp:362-400) of
 * [keyStart, keyEnd), host only.  cells_h[0..n] with cells_h[n] = keyEnd; returns n, -1 on error. */
int sphb200_span_sfc_range(uint64_t keyStart, uint64_t keyEnd, uint64_t* cells_h, size_t capacity);
/* cutSphere (grid.hpp:267-300): keep the particles with sqrt(x^2+y^2+z^2) <= radius, in place, order preserved */
int sphb200_cut_sphere(sphb200_ctx*, double radius, double* x, double* y, double* z, size_t n, uint64_t* count_h);
/* a1: MinMaxGpu / makeGlobalBox for open axes (primitives_gpu.h:52, box_mpi.hpp:84-119); minmax_h[6] host out */
int sphb200_minmax3(sphb200_ctx*, const double* x, const double* y, const double* z, size_t n, double* minmax_h);

/* ---- a3: sortByKeyGpu + sequenceGpu (primitives_gpu.h:87-111, primitives_gpu.cu:228-298) ----
 * keys sorted ascending in place (stable), perm[i] = original index of sorted element i */
int sphb200_sort_by_key(sphb200_ctx*, uint64_t* keys, uint32_t* perm, size_t n);
/* the same sort carrying the caller's values along (sortByKeyGpu(first, last, values, keyBuf, valueBuf),
 * primitives_gpu.h:96: the reference's GpuSfcSorter passes a sequence that starts at an offset); stable */
int sphb200_sort_pairs(sphb200_ctx*, uint64_t* keys, uint32_t* values, size_t n);

/* ---- a4: gatherGpu (primitives_gpu.h:62, primitives_gpu.cu:85-101), all fields in one launch ----
 * src_h/dst_h: host arrays of nfields device pointers; elem_bytes_h: 4 or 8 per field */
int sphb200_gather(sphb200_ctx*, const uint32_t* perm, size_t n, int nfields, const void* const* src_h,
                   void* const* dst_h, const int* elem_bytes_h);

/* ---- a5: computeNodeCountsGpu (csarray_gpu.h:56-67, csarray_gpu.cu:49-130) ---- */
int sphb200_node_counts(sphb200_ctx*, const uint64_t* leaves, int numLeaves, const uint64_t* keys, size_t n,
                        uint32_t* counts);

/* ---- a6/a7: updateOctreeGpu (tree/update_gpu.cuh:58-82) = rebalanceDecision + rebalanceTree + counts ----
 * leaves/counts are updated in place (capacity entries available); *numLeaves in/out; *converged out (host).
 * A leaf can split into up to 4096 (csarray.hpp:305): if the new tree does not fit, SPHB200_ERR_CAPACITY is returned
 * with *numLeaves_h = required number of leaves and leaves/counts untouched, so the call can be repeated. */
#define SPHB200_ERR_CAPACITY 2
int sphb200_update_octree(sphb200_ctx*, const uint64_t* keys, size_t n, unsigned bucketSize, uint64_t* leaves,
                          uint32_t* counts, int capacity, int* numLeaves_h, int* converged_h);
/* computeOctree (csarray.hpp:454-468): from {0, 2^63} until converged; *iterations_h = number of update steps */
int sphb200_compute_octree(sphb200_ctx*, const uint64_t* keys, size_t n, unsigned bucketSize, uint64_t* leaves,
                           uint32_t* counts, int capacity, int* numLeaves_h, int* iterations_h);
/* rebalanceDecisionGpu alone (csarray_gpu.h:70-78): ops[numLeaves] in {0,1,8,64,512,4096} */
int sphb200_rebalance_decision(sphb200_ctx*, const uint64_t* leaves, const uint32_t* counts, int numLeaves,
                               unsigned bucketSize, int32_t* ops, int* converged_h);

/* ---- a8: buildOctreeGpu (octree_gpu.h:50, octree_gpu.cu:55-175) ----
 * sizes: prefixes/internalToLeaf/leafToInternal [numNodes], childOffsets [numNodes+1], parents [(numNodes-1)/8],
 * levelRange [23]; numNodes = numLeaves + (numLeaves-1)/7 */
int sphb200_build_octree(sphb200_ctx*, const uint64_t* leaves, int numLeaves, uint64_t* prefixes, int32_t* childOffsets,
                         int32_t* parents, int32_t* levelRange, int32_t* internalToLeaf, int32_t* leafToInternal);
/* a9: computeGeoCentersGpu (source_center_gpu.h:50, source_center_gpu.cu:116-137) */
int sphb200_node_centers(sphb200_ctx*, const uint64_t* prefixes, int numNodes, const sphb200_box* box_h,
                         double* centers, double* sizes);
/* computeNodeLayout for one rank: exclusive scan of counts, layout[numLeaves+1] (domain/layout.hpp:150-163) */
int sphb200_layout(sphb200_ctx*, const uint32_t* counts, int numLeaves, uint32_t* layout);
/* The same scan for callers that must not allocate device memory (CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  No counterpart in the reference.  The tile
 * sums live in a file-scope `static __device__` array of the library (16 KB per GPU that loads it, held until exit)
 * instead of the context's grow-only arena; the call issues no cudaMalloc/cudaFree.  Limit: numLeaves <=
 * SPHB200_LAYOUT_NOALLOC_MAX (16 777 216 = 4096 tiles of 4096 counts), beyond it the call fails with an error message
 * and touches nothing.  All calls on one GPU share the array: calls that overlap on the device (two streams or two
 * contexts) race; calls on one stream are safe.  Same results as sphb200_layout, bit for bit. */
#define SPHB200_LAYOUT_NOALLOC_MAX 16777216
int sphb200_layout_noalloc(sphb200_ctx*, const uint32_t* counts, int numLeaves, uint32_t* layout);

/* ---- a13: neighbor search (traversal/find_neighbors.cuh:436-498 + xmass_gpu.cu:82-92 h iteration) ----
 * For every i in [first,last): all j != i with PBC-folded |r_ij|^2 < (2 h_i)^2 (findneighbors.hpp:95-165),
 * ascending j.  With iterate_h the coupled h/neighbor-count iteration of sph/find_neighbors.hpp:10-45 runs
 * (h updated in place, <= 10 iterations).  nc[i - first] = count + 1.
 * neighbors: opaque device buffer of sphb200_neighbors_size(last-first, ngmax) uint32 in the layout the pair
 *            operators read (lists interleaved per warp so that each warp-wide read is one 128-byte line; the layout
 *            is internal, convert with sphb200_neighbors_to_rowmajor / _from_rowmajor).
 * *numFails_h: non-converged particles. */
size_t sphb200_neighbors_size(size_t numTargets, unsigned ngmax);
int sphb200_find_neighbors(sphb200_ctx*, const sphb200_tree_view* tree_h, const double* x, const double* y,
                           const double* z, void* h, int dtype, uint32_t first, uint32_t last,
                           const sphb200_box* box_h, unsigned ng0, unsigned ngmax, int iterate_h, uint32_t* neighbors,
                           uint32_t* nc, unsigned* numFails_h);
/* internal layout <-> reference CPU layout neighbors[(i-first)*ngmax + k] (particles_data.hpp:235); nc[t] = count+1 */
int sphb200_neighbors_to_rowmajor(sphb200_ctx*, const uint32_t* neighbors, const uint32_t* nc, uint32_t numTargets,
                                  unsigned ngmax, uint32_t* rowmajor);
int sphb200_neighbors_from_rowmajor(sphb200_ctx*, const uint32_t* rowmajor, const uint32_t* nc, uint32_t numTargets,
                                    unsigned ngmax, uint32_t* neighbors);

/* ---- a14..a19: SPH-VE operators (sph/include/sph/sph_gpu.hpp:13-82) ----
 * nc is indexed like fields (nc = f->nc, entries [first,last) valid), neighbors as written by sphb200_find_neighbors for the same [first,last) */
int sphb200_xmass(sphb200_ctx*, int dtype, uint32_t first, uint32_t last, const sphb200_box* box_h,
                  const sphb200_hydro_const* hc_h, const uint32_t* neighbors, const sphb200_fields* f_h);
int sphb200_ve_def_gradh(sphb200_ctx*, int dtype, uint32_t first, uint32_t last, const sphb200_box* box_h,
                         const sphb200_hydro_const* hc_h, const uint32_t* neighbors, const sphb200_fields* f_h);
int sphb200_eos(sphb200_ctx*, int dtype, uint32_t first, uint32_t last, const sphb200_hydro_const* hc_h,
                const sphb200_fields* f_h);
int sphb200_iad_divv_curlv(sphb200_ctx*, int dtype, uint32_t first, uint32_t last, const sphb200_box* box_h,
                           const sphb200_hydro_const* hc_h, const uint32_t* neighbors, const sphb200_fields* f_h);
int sphb200_av_switches(sphb200_ctx*, int dtype, uint32_t first, uint32_t last, const sphb200_box* box_h,
                        const sphb200_hydro_const* hc_h, const uint32_t* neighbors, const sphb200_fields* f_h);
/* *minDtCourant_h (host out, may be NULL -> stays on device for sphb200_timestep) */
int sphb200_momentum_energy(sphb200_ctx*, int dtype, uint32_t first, uint32_t last, const sphb200_box* box_h,
                            const sphb200_hydro_const* hc_h, const uint32_t* neighbors, const sphb200_fields* f_h,
                            double* minDtCourant_h);

/* The same operator for the targets [begin,end) of [first,last) (begin - first a multiple of 128; lists and nc are
 * addressed relative to first like above).  pack != 0 rebuilds the gather records from the current field arrays;
 * pack == 0 reuses those of the previous call, so a caller may integrate [begin,end) of an earlier call in between
 * (sphb200_positions_update_h on that range): later ranges still see the pre-integration state of their neighbors.
 * Used with sphb200_av_switches_courant, which makes the time step known before this operator starts. */
int sphb200_momentum_energy_range(sphb200_ctx*, int dtype, uint32_t first, uint32_t last, uint32_t begin, uint32_t end,
                                  int pack, const sphb200_box* box_h, const sphb200_hydro_const* hc_h,
                                  const uint32_t* neighbors, const sphb200_fields* f_h);
/* hNew[i] = updateH(ng0, nc[i], h[i]) for [first,last) without touching h (update_h.hpp:12-24): the smoothing length
 * of the next step only depends on the neighbor counts, so a host-resident front-end can have it back early; nc, h,
 * hNew indexed like the fields */
int sphb200_update_h_into(sphb200_ctx*, int dtype, uint32_t first, uint32_t last, unsigned ng0, const uint32_t* nc,
                          const void* h, void* hNew);
/* sphb200_av_switches that also evaluates the Courant step of momentumEnergy (same vijsignal = 0.5 (c_i + c_j) - 2 w_ij
 * over the same pairs, bit-identical to what sphb200_momentum_energy finds) and leaves it on the device for
 * sphb200_timestep.  Not available with avClean (there w_ij needs the velocity-gradient correction). */
int sphb200_av_switches_courant(sphb200_ctx*, int dtype, uint32_t first, uint32_t last, const sphb200_box* box_h,
                                const sphb200_hydro_const* hc_h, const uint32_t* neighbors, const sphb200_fields* f_h);

/* ---- a20: rhoTimestep + computeTimestep local part (sph/include/sph/ts_global.hpp:71-112) ----
 * out_h[0] = minDtCourant (from the last momentum_energy call), out_h[1] = Krho/|max divv|,
 * out_h[2] = min(out[0], out[1], 1.1*minDt)  (the value fed to the global MIN all-reduce) */
int sphb200_timestep(sphb200_ctx*, int dtype, uint32_t first, uint32_t last, const sphb200_hydro_const* hc_h,
                     const sphb200_fields* f_h, double* out_h);
/* device address of the 3 doubles above, for an in-stream ncclAllReduce(min) */
double* sphb200_timestep_device_ptr(sphb200_ctx*);

/* ---- turbulence stirring (sph/include/sph/hydro_turb/*, main/src/propagator/turb_ve.hpp:113-118) ----
 * Operator level: computeStirringGpu (stirring_gpu.cu:43-93).  modes[numModes][3], phaseReal/phaseImag[numModes][3],
 * amplitudes[numModes] are device arrays of double; ax, ay, az of `real` type get solWeightNorm * sum over modes added. */
int sphb200_stirring(sphb200_ctx*, int dtype, uint32_t first, uint32_t last, const double* x, const double* y,
                     const double* z, void* ax, void* ay, void* az, int numModes, const double* modes,
                     const double* phaseReal, const double* phaseImag, const double* amplitudes, double solWeightNorm);

/* Driver level: sph::TurbulenceData + driveTurbulence (turbulence_data.hpp:57-183, driver.hpp:79-130).  The mode
 * table, the Ornstein-Uhlenbeck sequence (std::mt19937 / std::normal_distribution, same stream as the reference when
 * built against the same C++ library) and the phase projection live on the host. */
typedef struct sphb200_turb_config
{
    double solWeight, Lbox, stEnergyPrefac, stMachVelocity, epsilon, powerLawExp, anglesExp;
    uint64_t stMaxModes, rngSeed;
    int stSpectForm; /* 0 band, 1 parabola, 2 power law */
} sphb200_turb_config;
typedef struct sphb200_turb sphb200_turb;
/* host work only (mode table, initial OU state): no device needed; the device copies are made by the first stir */
int sphb200_turb_create(const sphb200_turb_config*, sphb200_turb** out);
void sphb200_turb_destroy(sphb200_turb*);
int sphb200_turb_num_modes(const sphb200_turb*);
/* host copies; any pointer may be NULL.  scalars4 = {variance, decayTime, solWeight, solWeightNorm};
 * modes[3n], amplitudes[n], phases[6n] (OU state), phasesReal[3n], phasesImag[3n] */
int sphb200_turb_get(const sphb200_turb*, double* scalars4, double* modes, double* amplitudes, double* phases,
                     double* phasesReal, double* phasesImag);
/* restart (TurbulenceData::loadOrStore): OU state and/or the text form of the random engine */
int sphb200_turb_set_state(sphb200_turb*, const double* phases, const char* rngState);
int sphb200_turb_rng_state(const sphb200_turb*, char* buf, size_t bufSize); /* returns the size needed */
/* the two halves of driveTurbulence, for callers that stir range by range (sphb200_sim_step_host, the multi-rank
 * host): advance = OU step by dt + projection + upload of the phases (once per step), stir = the kernel on [first,last).
 * advance with ctx == NULL is the host half alone (the next stir uploads the phases) */
int sphb200_turb_advance(sphb200_turb*, sphb200_ctx*, double dt);
int sphb200_turb_stir(sphb200_turb*, sphb200_ctx*, int dtype, uint32_t first, uint32_t last, const double* x,
                      const double* y, const double* z, void* ax, void* ay, void* az);
/* driveTurbulence: advance the OU phases by dt, project, upload, add the stirring accelerations on [first,last) */
int sphb200_turb_drive(sphb200_turb*, sphb200_ctx*, int dtype, uint32_t first, uint32_t last, const double* x,
                       const double* y, const double* z, void* ax, void* ay, void* az, double dt);

/* ---- observables: computeConservedQuantities (main/src/observables/conserved_quantities.hpp:50-180) ----
 * out[10] = { eKin = 1/2 sum m v^2, eInt = sum cv T m, egrav (0), linmom[3], angmom[3], sum of nc } over [first,last),
 * accumulated in double in a fixed order (reproducible).  out_h may be NULL: the values stay on the device
 * (sphb200_conserved_device_ptr) for an in-stream all-reduce over ranks. */
int sphb200_conserved_quantities(sphb200_ctx*, int dtype, uint32_t first, uint32_t last,
                                 const sphb200_hydro_const* hc_h, const sphb200_fields* f_h, double* out_h);
double* sphb200_conserved_device_ptr(sphb200_ctx*);
/* localMachSquareSum / machSquareSumGpu (main/src/observables/turbulence_mach_rms.hpp:38-56, gpu_reductions.cu:110-122):
 * sum over [first,last) of |v|^2 / c^2 (c as left by sphb200_eos); the RMS Mach number of TurbulenceMachRMS is
 * sqrt(sum over ranks / numParticlesGlobal) (:70-83).  Fixed-order fp64 reduction; the sum also stays on the device
 * (sphb200_mach_device_ptr) for an in-stream all-reduce. */
int sphb200_mach_square_sum(sphb200_ctx*, int dtype, uint32_t first, uint32_t last, const sphb200_fields* f_h,
                            double* out_h);
double* sphb200_mach_device_ptr(sphb200_ctx*);

/* The pair operators gather neighbor data from packed copies ("gather records") of the field arrays kept in the
 * context.  Stand-alone operator calls repack on every call; a sphb200_sim reuses them across operators and
 * invalidates them itself.  Call this after writing to a sim's device arrays obtained through sphb200_sim_fields()
 * at a later time than the getter call itself. */
void sphb200_invalidate_records(sphb200_ctx*);

/* The pair operators that interpolate the kernel table once or twice per pair (XMass, IadDivvCurlv, AVswitches,
 * MomentumEnergy) exist in two variants that return the same bits: the table read through L1 (fastest on lattices)
 * and the table staged in shared memory of a persistent CTA (fastest on disordered particle sets).  By default the
 * context times both on the caller's data and keeps the faster one (variant -1); 0 / 1 pin the L1 / shared-memory
 * variant (tests, profiling; also SPHB200_TABLE=l1|smem in the environment).  No counterpart in the reference, whose
 * kernels read the table from global memory (sph/table_lookup.hpp:13-25). */
int sphb200_set_table_variant(sphb200_ctx*, int variant);

/* ---- a21: computePositionsGpu + updateSmoothingLengthGpu (positions_gpu.cu:119-192, update_h_gpu.cu:40-62) ---- */
int sphb200_positions_update_h(sphb200_ctx*, int dtype, uint32_t first, uint32_t last, const sphb200_box* box_h,
                               const sphb200_hydro_const* hc_h, double dt, double dt_m1, const sphb200_fields* f_h);

/* ---- halo pack / unpack: gatherRanges / scatterRanges (halos/gather_halos_gpu.h:38-46) ----
 * copies numRanges index ranges [rangeBegin[r], rangeEnd[r]) of nfields arrays into/out of one packed buffer
 * (field-major, ranges concatenated).  range arrays are device pointers. */
int sphb200_pack_ranges(sphb200_ctx*, const uint32_t* rangeScan, const uint32_t* rangeBegin, int numRanges,
                        uint32_t totalElems, int nfields, const void* const* src_h, const int* elem_bytes_h,
                        void* packed);
int sphb200_unpack_ranges(sphb200_ctx*, const uint32_t* rangeScan, const uint32_t* rangeBegin, int numRanges,
                          uint32_t totalElems, int nfields, void* const* dst_h, const int* elem_bytes_h,
                          const void* packed);
/* All peers of one halo exchange in one launch.  Ranges are concatenated in peer order (peerCount_h[q] elements
 * belong to peer q); the packed buffer holds one block per peer, inside a block the fields follow one another, each
 * 8-byte aligned -- exactly the split sizes of an all-to-all (returned in bytesPerPeer_h, may be NULL; packed may be
 * NULL to query the sizes only). */
int sphb200_pack_ranges_peers(sphb200_ctx*, const uint32_t* rangeScan, const uint32_t* rangeBegin, int numRanges,
                              int numPeers, const uint32_t* peerCount_h, int nfields, const void* const* src_h,
                              const int* elem_bytes_h, void* packed, size_t* bytesPerPeer_h);
int sphb200_unpack_ranges_peers(sphb200_ctx*, const uint32_t* rangeScan, const uint32_t* rangeBegin, int numRanges,
                                int numPeers, const uint32_t* peerCount_h, int nfields, void* const* dst_h,
                                const int* elem_bytes_h, const void* packed, size_t* bytesPerPeer_h);

/* ---- halo discovery (multi-GPU) ----
 * segmentMaxGpu (primitives_gpu.h:76, primitives_gpu.cu:180-207): out[s] = max h over [layout[s], layout[s+1]) */
int sphb200_segment_max(sphb200_ctx*, const void* h, int dtype, const uint32_t* layout, int numSegments, double* out);
/* owner-side findHalos (traversal/collisions_gpu.h:59, collisions_gpu.cu:39-66): mask[leaf] bit q set if the leaf box
 * (center/half-size, [numLeaves][3]) overlaps one of peer q's query boxes; boxes: [numPeers][maxBoxes][6]
 * (center xyz, half-size xyz), boxCount[numPeers] (device) */
int sphb200_mark_halo_leaves(sphb200_ctx*, const double* leafCenters, const double* leafSizes, int numLeaves,
                             const double* boxes, const int* boxCount, int numPeers, int maxBoxes, int myRank,
                             const sphb200_box* box_h, uint32_t* mask);

/* ======================= tier 2: domain + propagator ======================= */
/* cstone::Domain<uint64_t,double,GpuTag> (domain/include/cstone/domain/domain.hpp:65-683) and
 * sphexa::HydroVeProp (main/src/propagator/ve_hydro.hpp:54-219) for one rank of an SFC-partitioned run.
 * The simulation object owns all device arrays ([halos|assigned|halos] layout, SFC order). */
typedef struct sphb200_sim sphb200_sim;

typedef struct sphb200_sim_config
{
    int dtype;            /* SPHB200_F64 | SPHB200_MIXED */
    int rank, numRanks;   /* SFC partition; numRanks==1 -> no halos */
    unsigned bucketSize;      /* global tree bucket, sphexa.cpp:133-135: max(64, N/(100 P)) */
    unsigned bucketSizeFocus; /* 64 */
    size_t capacity;      /* max particles incl. halos on this rank */
    sphb200_box box;
    sphb200_hydro_const hc; /* wh/whd are HOST pointers here (copied to device) */
} sphb200_sim_config;

/* host field table for upload/download: same names as sphb200_fields, host pointers; NULL = skip */
int sphb200_sim_create(sphb200_ctx*, const sphb200_sim_config* cfg, sphb200_sim** out);
void sphb200_sim_destroy(sphb200_sim*);
/* upload n particles (conserved fields) into the assigned range; any order */
int sphb200_sim_upload(sphb200_sim*, size_t n, const sphb200_fields* host_fields);
/* download assigned range [startIndex,endIndex) of every non-NULL field */
int sphb200_sim_download(sphb200_sim*, const sphb200_fields* host_fields, uint64_t* keys_h);
/* TurbVeProp: stir after the hydro forces of every step with dt = minDt (turb_ve.hpp:113-118); NULL detaches.
 * The sim does not own the turbulence object. */
int sphb200_sim_attach_turbulence(sphb200_sim*, sphb200_turb*);
/* computeConservedQuantities for the owned particles (see sphb200_conserved_quantities) */
int sphb200_sim_conserved(sphb200_sim*, double* out10_h);
/* TurbulenceMachRMS::computeAndWrite for one rank: sqrt(machSquareSum / numParticlesGlobal); numParticlesGlobal = 0:
 * the owned count */
int sphb200_sim_mach_rms(sphb200_sim*, uint64_t numParticlesGlobal, double* out_h);
/* One time step on host-resident state: same results as upload + step + download, with the PCIe transfers staged
 * beside the kernels ({x,y,z} first, {h,m,temp} next, the rest behind the neighbor search; m and alpha return while
 * MomentumEnergy runs).  host_in/host_out: all 14 conserved fields, pinned memory for overlap.  Single rank only. */
int sphb200_sim_step_host(sphb200_sim*, size_t n, const sphb200_fields* host_in, const sphb200_fields* host_out);
/* Domain::sync (domain.hpp:196-244): keys, sort, gather, tree update, layout.  Multi-rank pieces
 * (assignment, migration, halo discovery) are driven by the host through the calls further down. */
int sphb200_sim_sync(sphb200_sim*);
/* HydroVeProp::computeForces without the sync (ve_hydro.hpp:140-192), single rank */
int sphb200_sim_compute_forces(sphb200_sim*);
/* HydroVeProp::integrate (ve_hydro.hpp:207-217): computeTimestep + computePositions + updateSmoothingLength;
 * globalMinDt < 0: use the local minimum */
int sphb200_sim_integrate(sphb200_sim*, double globalMinDt);
/* one full step = sync + computeForces + integrate (single rank) */
int sphb200_sim_step(sphb200_sim*);
/* step phases for multi-rank drivers: phase k runs the operators between halo exchanges k and k+1 of
 * ve_hydro.hpp:149-192: 0: neighbors+XMass | 1: VeDefGradh+EOS | 2: IadDivvCurlv | 3: AVswitches | 4: MomentumEnergy */
int sphb200_sim_forces_phase(sphb200_sim*, int phase);

/* getters (domain.hpp:382-429) */
uint32_t sphb200_sim_start_index(const sphb200_sim*);
uint32_t sphb200_sim_end_index(const sphb200_sim*);
uint32_t sphb200_sim_n_with_halos(const sphb200_sim*);
int sphb200_sim_num_leaves(const sphb200_sim*);
int sphb200_sim_num_global_leaves(const sphb200_sim*);
double sphb200_sim_min_dt(const sphb200_sim*);
double sphb200_sim_ttot(const sphb200_sim*);
uint64_t sphb200_sim_total_neighbors(sphb200_sim*); /* sum of nc over assigned, the "Total Neighbors" check line */
/* device field table of the simulation (for halo exchange / tests) */
int sphb200_sim_fields(sphb200_sim*, sphb200_fields* out);
int sphb200_sim_tree(sphb200_sim*, sphb200_tree_view* out);
const uint64_t* sphb200_sim_keys(const sphb200_sim*);
const uint32_t* sphb200_sim_neighbors(const sphb200_sim*);
/* per-phase device times of the last step in ms: sync, neighbors, xmass, gradh, eos, iad, av, momentum, timestep+update */
int sphb200_sim_phase_times(sphb200_sim*, float* ms_out9);
/* local (pre-all-reduce) time step of the last computeForces: out3 = {minDtCourant, minDtRho, min incl. 1.1*minDt} */
int sphb200_sim_local_timestep(sphb200_sim*, double* out3_h);
/* device views of the trees: replicated global tree (assignment.hpp:62-113) and focus-tree leaf counts */
const uint64_t* sphb200_sim_global_leaves(const sphb200_sim*);
const uint32_t* sphb200_sim_global_counts(const sphb200_sim*);
const uint32_t* sphb200_sim_leaf_counts(const sphb200_sim*);
void sphb200_sim_get_box(const sphb200_sim*, sphb200_box* out);
/* override minDt / minDt_m1 (restart, tests) */
void sphb200_sim_set_dt(sphb200_sim*, double minDt, double minDt_m1);
/* set layout for multi-rank: arrays hold n particles, assigned = [start,end) */
int sphb200_sim_set_layout(sphb200_sim*, uint32_t start, uint32_t end, uint32_t nWithHalos);

/* ======================= multi-rank domain (SFC-partitioned run over several GPUs) =======================
 * cstone::Domain<uint64_t,double,GpuTag> for P ranks (domain/include/cstone/domain/domain.hpp:65-683) with the
 * collectives issued from C++ on the context's stream.  Two transports:
 *  - NCCL (one process per GPU): rank 0 calls sphb200_comm_unique_id, the front-end broadcasts the 128 bytes
 *    (MPI_Bcast in the reference's main), every rank calls sphb200_comm_create_nccl;
 *  - loopback: P virtual ranks inside ONE process on ONE GPU, exchanges are device-to-device copies.  It exists so
 *    that every multi-rank code path can be tested on a single GPU.
 * Collective entry points take the handles of ALL ranks the calling process drives: one with NCCL, all P in rank
 * order with the loopback (all on one stream). */
typedef struct sphb200_comm sphb200_comm;
int sphb200_comm_create_loopback(int numRanks, sphb200_comm** out);
int sphb200_comm_unique_id(void* id128);
int sphb200_comm_create_nccl(int device, int rank, int numRanks, const void* id128, sphb200_comm** out);
void sphb200_comm_destroy(sphb200_comm*);
int sphb200_comm_num_ranks(const sphb200_comm*);
int sphb200_comm_is_loopback(const sphb200_comm*);

/* Domain(rank, nRanks, bucketSize, bucketSizeFocus, theta, box) (domain.hpp:97-113): cfg.rank / cfg.numRanks name the
 * rank; the handle is a sphb200_sim (it owns the particle arrays), so upload / download / getters / sphb200_sim_fields
 * work on it.  The communicator must outlive the handle. */
int sphb200_domain_create(sphb200_ctx*, const sphb200_sim_config* cfg, sphb200_comm* comm, sphb200_sim** out);
/* Domain::sync (domain.hpp:196-244): global box, keys, sort, replicated global tree (counts all-reduced,
 * tree/update_mpi.hpp:81-104), assignment with limitBoundaryShifts (domaindecomp.hpp:115-181), migration
 * (domaindecomp_mpi_gpu.cuh:85-186), focus tree, halo discovery (findHalos criterion, traversal/collisions.hpp:78-105,
 * evaluated by the owner), layout [halos | assigned | halos], exchange of x, y, z, h, m, search tree */
int sphb200_domain_sync(sphb200_sim* const* local, int numLocal);
/* Domain::exchangeHalos (domain.hpp:375-379, halos/exchange_halos_gpu.cuh:50-143) for the given fields;
 * field id = position of the member in sphb200_fields (0 = x, 1 = y, ... 14 = xm, ... 31 = nc) */
int sphb200_domain_exchange_halos(sphb200_sim* const* local, int numLocal, const int* fieldIds, int numFields);
/* HydroVeProp on P ranks (main/src/propagator/ve_hydro.hpp:132-217): sync + computeForces with the four halo
 * exchanges between the operators; integrate with the MIN all-reduce of the time step (ts_global.hpp:104-106) */
int sphb200_domain_compute_forces(sphb200_sim* const* local, int numLocal);
int sphb200_domain_integrate(sphb200_sim* const* local, int numLocal);
int sphb200_domain_step(sphb200_sim* const* local, int numLocal);
/* one step on host-resident state (the call a host-resident front-end makes): upload of the owned conserved fields
 * from pinned memory, sync + forces + integrate, download of the owned range after the step.  n_in / host_in / host_out
 * / n_out: one entry per local rank; host_out arrays hold hostCapacity elements each (ownership moves between ranks) */
int sphb200_domain_step_host(sphb200_sim* const* local, int numLocal, const size_t* n_in, const sphb200_fields* host_in,
                             const sphb200_fields* host_out, size_t* n_out, size_t hostCapacity);
/* computeConservedQuantities summed over the ranks (conserved_quantities.hpp:113-180) */
int sphb200_domain_conserved(sphb200_sim* const* local, int numLocal, double* out10_h);
/* Domain::setHaloFactor (domain.hpp): widens the halo search radius (block time steps) */
void sphb200_domain_set_halo_factor(sphb200_sim*, float factor);
typedef struct sphb200_domain_info_t
{
    int rank, numRanks;
    uint32_t startIndex, endIndex, nWithHalos;
    int numGlobalLeaves, numFocusLeaves, numSearchLeaves;
    uint32_t numSurfaceBoxes; /* halo boxes this rank published in the last sync */
    uint32_t numSendRanges;
    uint32_t migratedIn, migratedOut;
    uint64_t haloBytes;       /* bytes sent in halo exchanges since creation */
    uint32_t haloSend[32], haloRecv[32]; /* particles per peer */
    uint64_t assignment[33];  /* SfcAssignment: rank r owns keys [assignment[r], assignment[r+1]) */
    /* device time of the halo exchanges of the last sync + compute_forces (pack, transfer, unpack; CUDA events on
     * the rank's stream): [0] x,y,z,h,m inside the sync, [1..4] the exchanges before gradh, IAD, AVswitches and
     * MomentumEnergy (ve_hydro.hpp:149-192), [5] their sum; 0 before the first step */
    float haloExchangeMs[6];
} sphb200_domain_info_t;
int sphb200_domain_info(const sphb200_sim*, sphb200_domain_info_t* out);

/* host-only pieces of the assignment (no device needed):
 * makeSfcAssignment + limitBoundaryShifts (domaindecomp.hpp:115-181); oldBounds_h may be NULL (first call) */
int sphb200_sfc_assignment(int numRanks, const uint32_t* counts_h, const uint64_t* leaves_h, int numLeaves,
                           const uint64_t* oldBounds_h, uint64_t* bounds_h, uint32_t* rankCounts_h);
/* the tree GlobalAssignment starts from: computeSpanningTree(initialDomainSplits(P, log8ceil(100 P)))
 * (assignment.hpp:62-69); returns the number of leaves (leaves_h gets one more key), -needed if capacity is too small */
int sphb200_initial_global_tree(int numRanks, uint64_t* leaves_h, int capacity);
/* findHalosGpu (traversal/collisions_gpu.h:59, collisions.hpp:78-105) on one linked octree: flags[leaf] = 1 for every
 * leaf outside [firstLeaf, lastLeaf) whose box overlaps the halo box of a leaf inside; radii: float per leaf;
 * flags zeroed by the caller; *numBoxes_h (may be NULL) = leaves whose halo box leaves the range */
int sphb200_find_halos(sphb200_ctx*, const uint64_t* prefixes, const int32_t* childOffsets, const int32_t* internalToLeaf,
                       const uint64_t* leaves, const float* radii, const sphb200_box* box_h, int firstLeaf, int lastLeaf,
                       int* flags, int* numBoxes_h);
/* per-leaf halo radius as Halos::discover computes it (halos/halos.hpp:166-188): float(max h of the leaf * 2 *
 * searchExt), 0 for empty leaves */
int sphb200_halo_radii(sphb200_ctx*, const void* h, int dtype, const uint32_t* layout, int firstLeaf, int lastLeaf,
                       float searchExt, float* radii);

/* ---- a10: the locally essential ("focused") octree: particle counts decide inside the focus key range, a
 * minimum-distance multipole acceptance criterion (MAC) decides outside (csrc/focus.cu) ----
 * Tier 1, device arrays of a linked octree as sphb200_build_octree makes them (numNodes = leaves + internal nodes):
 *
 * computeMinMacR2 (traversal/macs.hpp:58-71) per node: centers4[4i..4i+2] = geometric center, [4i+3] = squared
 * acceptance radius (2 max(size) invThetaEff)^2; invThetaEff = 1/theta + 0.5 (invThetaMinMac, macs.hpp:44) */
int sphb200_min_mac_r2(sphb200_ctx*, const uint64_t* prefixes, int numNodes, float invThetaEff, const sphb200_box* box_h,
                       double* centers4);
/* markMacs / markMacsGpu (traversal/macs.hpp:240-264, macs_gpu.cu): markings[i] = 1 iff node i fails evaluateMacPbc
 * paired with a cell of [focusStart, focusEnd) -- for one spanSfcRange cell of the focus the node and all its
 * ancestors violate the MAC and none of them lies inside the focus; all other entries are set to 0.  Takes the
 * `parents` array where the reference's traversal takes childOffsets. */
int sphb200_mark_macs(sphb200_ctx*, const uint64_t* prefixes, const int32_t* parents, const double* centers4, int numNodes,
                      const sphb200_box* box_h, uint64_t focusStart, uint64_t focusEnd, char* markings);
/* rebalanceDecisionEssential[Gpu] (focus/rebalance.hpp:144-160, mergeCountAndMacOp :49-90): nodeOps[i] = 0 merge,
 * 1 keep, 8 split, for ALL nodes; counts/macs per node */
int sphb200_rebalance_decision_essential(sphb200_ctx*, const uint64_t* prefixes, const int32_t* childOffsets,
                                         const int32_t* parents, const uint32_t* counts, const char* macs,
                                         uint64_t focusStart, uint64_t focusEnd, unsigned bucketSize, int numNodes,
                                         int32_t* nodeOps);
/* enforceKeys[Gpu] (focus/rebalance.hpp:189-258): keys (device) that the next tree must contain; cancels merges of
 * their supporting ancestors and splits the closest leaf by one level.  *status_h = ResolutionStatus: 0 converged,
 * 1 cancelMerge, 2 rebalance, 3 failed (a key lies more than one level below the closest leaf) */
int sphb200_enforce_keys(sphb200_ctx*, const uint64_t* keys, int numKeys, const uint64_t* prefixes,
                         const int32_t* childOffsets, const int32_t* parents, int32_t* nodeOps, int* status_h);
/* protectAncestors[Gpu] (focus/rebalance.hpp:103-176): a merged node that is the left-most descendant of its closest
 * non-zero ancestor takes that ancestor's op; *converged_h = 1 if every node ends with op 1 */
int sphb200_protect_ancestors(sphb200_ctx*, const uint64_t* prefixes, const int32_t* parents, int numNodes,
                              int32_t* nodeOps, int* converged_h);
/* findPeersMac (traversal/peers.hpp:63-120), host only: peers_h[r] = 1 for every rank that owns a leaf of the
 * replicated global tree which fails the mutual min-distance MAC paired with a leaf of myRank's segment (dual
 * traversal, the reference's split rule); bounds_h = SfcAssignment, numRanks + 1 keys that are leaf keys */
int sphb200_find_peers_mac(int myRank, int numRanks, const uint64_t* bounds_h, const uint64_t* leaves_h, int numLeaves,
                           const sphb200_box* box_h, float invThetaEff, int* peers_h);

/* Tier 2: the tree object.  sphb200_focus_update == FocusedOctreeSingleNode::update (focus/octree_focus.hpp:236-266)
 * = CombinedUpdate::updateFocusGpu (:158-213: decision, enforceKeys on {focusStart, focusEnd, mandatoryKeys},
 * protectAncestors, rebalanceTree, injectKeys when enforcing failed, linked octree) with the counts and MACs of the
 * previous call, then MACs (computeMinMacR2 + markMacs) and counts of the new tree from `keys` (sorted, device: ALL
 * particles the counts shall see).  The three stages are also callable one by one in the order of
 * FocusedOctree::converge (octree_focus_mpi.hpp:498-516: updateMinMac, updateTree, updateCounts), with
 * sphb200_focus_set_leaf_counts for counts that were reduced over ranks.  *converged_h = 1 if the tree structure did
 * not change.  Starts as the root node with count bucketSize + 1 and MAC 1, like the reference. */
typedef struct sphb200_focus sphb200_focus;
typedef struct sphb200_focus_view
{
    int numLeaves, numNodes;
    int lastStatus;             /* ResolutionStatus of the last update_tree */
    const uint64_t* leaves;     /* numLeaves + 1 */
    const uint32_t* leafCounts; /* numLeaves */
    const uint64_t* prefixes;   /* linked octree, numNodes */
    const int32_t *childOffsets, *parents, *levelRange, *internalToLeaf, *leafToInternal;
    const uint32_t* counts; /* numNodes, NULL while stale */
    const char* macs;       /* numNodes, NULL while stale */
    const double* centers;  /* 4 x numNodes, NULL while stale */
} sphb200_focus_view;
int sphb200_focus_create(sphb200_ctx*, unsigned bucketSize, float theta, sphb200_focus** out);
void sphb200_focus_destroy(sphb200_focus*);
int sphb200_focus_update(sphb200_focus*, const sphb200_box* box_h, const uint64_t* keys, size_t n, uint64_t focusStart,
                         uint64_t focusEnd, const uint64_t* mandatoryKeys_h, int numMandatory, int* converged_h);
int sphb200_focus_update_tree(sphb200_focus*, uint64_t focusStart, uint64_t focusEnd, const uint64_t* mandatoryKeys_h,
                              int numMandatory, int* converged_h);
int sphb200_focus_update_min_mac(sphb200_focus*, const sphb200_box* box_h, uint64_t focusStart, uint64_t focusEnd);
int sphb200_focus_update_counts(sphb200_focus*, const uint64_t* keys, size_t n);
int sphb200_focus_set_leaf_counts(sphb200_focus*, const uint32_t* leafCounts);
int sphb200_focus_get_view(sphb200_focus*, sphb200_focus_view* out);
/* The same tree for every rank of a multi-rank domain, on request (the step path does not need it: halo discovery is
 * owner-side on the guided tree).  One call = one round of what Domain::sync does with its FocusedOctree
 * (domain.hpp:214-227): findPeersMac on the replicated global tree, updateMinMac, updateTree with the peers' boundaries
 * enforced, updateCounts with GLOBAL counts (one all-gather of the ranks' leaf arrays, each rank counts its own
 * particles in all of them, one all-reduce -- in place of the treelet exchange, exchange_focus.hpp:113-198); the first
 * call iterates until every rank has converged (FocusedOctree::converge, octree_focus_mpi.hpp:498-516).  Collective:
 * pass the handles as for sphb200_domain_sync, after a sync.  *converged_h = 1 if no rank's tree changed.
 * sphb200_domain_focus_view: the tree of one rank (device pointers, valid until the next update) and, if peers_h is
 * not NULL, the numRanks flags of findPeersMac.  focusTransfer (exchange_focus.hpp:219-290) is not carried out. */
int sphb200_domain_update_focus(sphb200_sim* const* sims, int numLocal, float theta, int* converged_h);
int sphb200_domain_focus_view(const sphb200_sim*, sphb200_focus_view* out, int* peers_h);

/* ---- f4: restart files ----
 * The reference checkpoints through H5Part/HDF5 (main/src/io/ifile_io_hdf5.cpp, sphexa.cpp:232-246): per dump one
 * "Step#k" group with the step attributes of ParticlesData::loadOrStoreAttributes (sph/particles_data.hpp:142-193),
 * Box::loadOrStore (cstone/sfc/box.hpp:168-175), TurbulenceData::loadOrStore (hydro_turb/turbulence_data.hpp:90-125)
 * and one dataset per conserved field (Propagator::conservedFields).  HDF5 is not available on this box, so the same
 * content goes into a self-describing little-endian binary file: a directory of named attributes followed by named
 * SoA field blocks, one file per rank.  Attribute and field names are the reference's.  A restarted run continues
 * bit-identically (tests/test_gpu_restart.py). */
int sphb200_sim_get_config(const sphb200_sim*, sphb200_sim_config* out); /* hc.wh/whd: device pointers */
/* replace the scalar constants (Kcour, Krho, gamma, muiConst, alpha*, decay_constant, ng0, At*, ramp);
 * ngmax, K, avClean and the kernel tables are fixed at creation and must agree */
int sphb200_sim_set_constants(sphb200_sim*, const sphb200_hydro_const*);
void sphb200_sim_set_time(sphb200_sim*, double ttot);
void sphb200_sim_set_box(sphb200_sim*, const sphb200_box*);
sphb200_turb* sphb200_sim_turbulence(const sphb200_sim*);

typedef struct sphb200_restart sphb200_restart;
typedef struct sphb200_restart_info
{
    uint64_t iteration, numParticlesGlobal, numParticles; /* numParticles: in this file */
    int dtype;                                            /* precision the fields were written in */
    int hasTurbulence;
    double time, minDt, minDt_m1;
    sphb200_box box;
} sphb200_restart_info;
/* the writer underneath sphb200_sim_save, for drivers that keep their particle state outside a sphb200_sim (the
 * multi-rank host): host_fields = the 14 conserved fields of info->numParticles particles in precision info->dtype;
 * turb may be NULL */
int sphb200_restart_write(const char* path, const sphb200_restart_info* info, const sphb200_hydro_const* hc,
                          const sphb200_turb* turb, int rank, int numRanks, const sphb200_fields* host_fields);
/* block-time-step state of a dump: HydroVeBdtProp::save writes Timestep::loadOrStore with the prefix "ts::"
 * (ve_hydro_bdt.hpp:153, sph/include/sph/timestep.h:29-33) and "rung" is one of its conserved fields
 * (ve_hydro_bdt.hpp:94).  rung: one byte per particle of the file, host memory. */
typedef struct sphb200_restart_bdt
{
    int numRungs;
    float dt_m1[4];
    const uint8_t* rung;
} sphb200_restart_bdt;
/* sphb200_restart_write plus the attributes ts::numRungs, ts::dt_m1 and the particle field "rung"; bdt may be NULL */
int sphb200_restart_write_bdt(const char* path, const sphb200_restart_info* info, const sphb200_hydro_const* hc,
                              const sphb200_turb* turb, int rank, int numRanks, const sphb200_fields* host_fields,
                              const sphb200_restart_bdt* bdt);
/* write the owned particles [startIndex,endIndex): attributes + the 14 conserved fields (+ turbulence state when a
 * driver is attached, + the block-time-step state after sphb200_sim_bdt_enable).  iteration / numParticlesGlobal are
 * the caller's (they live in the reference's main loop). */
int sphb200_sim_save(sphb200_sim*, const char* path, uint64_t iteration, uint64_t numParticlesGlobal);

int sphb200_restart_open(const char* path, sphb200_restart** out);
void sphb200_restart_close(sphb200_restart*);
int sphb200_restart_get_info(const sphb200_restart*, sphb200_restart_info* out);
/* generic attribute access (H5Part stepAttribute): returns the element count, -1 if absent; copies at most
 * `capacity` elements of the stored type (8-byte double/int64 or bytes) */
int64_t sphb200_restart_attribute(const sphb200_restart*, const char* name, void* out, size_t capacity);
/* read particles [first, first+count) of every field with a non-NULL host pointer, converted to `dtype` when the
 * file holds the other precision (fileutils readField semantics); a rank assembles its slice from several files by
 * offsetting the pointers */
int sphb200_restart_read_fields(const sphb200_restart*, uint64_t first, uint64_t count, int dtype,
                                const sphb200_fields* host_fields);
/* particles [first, first+count) of a one-byte-per-particle field ("rung"); an error when the file has none */
int sphb200_restart_read_bytes(const sphb200_restart*, const char* name, uint64_t first, uint64_t count, uint8_t* out);
/* time, minDt, minDt_m1, the optional constants found in the file, the box, the turbulence state (when the file
 * has one and a driver is attached) and ts::numRungs / ts::dt_m1 (when the file has them and block time steps are
 * enabled on the sim: HydroVeBdtProp::load, ve_hydro_bdt.hpp:155-168).  The "rung" field travels with the particle
 * slices: sphb200_restart_read_bytes + sphb200_sim_bdt_upload_rungs. */
int sphb200_restart_apply(const sphb200_restart*, sphb200_sim*);

/* ---- f3: block time steps (HydroVeBdtProp, main/src/propagator/ve_hydro_bdt.hpp) ----
 * cstone::GroupView (domain/include/cstone/traversal/groups.hpp): numGroups runs [groupStart[g], groupEnd[g]) of
 * SFC-consecutive particles inside [firstBody, lastBody); device arrays */
typedef struct sphb200_group_view
{
    uint32_t firstBody, lastBody, numGroups;
    const uint32_t* groupStart;
    const uint32_t* groupEnd;
} sphb200_group_view;
/* computeSpatialGroups (sph/include/sph/groups.cu:30-47 -> cstone::computeGroupSplits, traversal/groups.cuh:266-330):
 * runs of 32 from `first`, cut where consecutive particles are further apart than tolFactor x the smallest leaf edge
 * under the run.  groups[0..numGroups] = delimiters (groupStart = groups, groupEnd = groups + 1); capacity in groups;
 * SPHB200_ERR_CAPACITY with the required count in *numGroups_h when it does not fit */
int sphb200_compute_groups(sphb200_ctx*, uint32_t first, uint32_t last, const double* x, const double* y, const double* z,
                           const uint64_t* leaves, int numLeaves, const uint32_t* layout, const sphb200_box* box_h,
                           float tolFactor, uint32_t* groups, uint32_t capacity, uint32_t* numGroups_h);
/* groupDivvTimestepGpu / groupAccTimestepGpu (sph/include/sph/ts_groups.cu:16-78): groupDt[g] = min(groupDt[g], ...) */
int sphb200_group_divv_timestep(sphb200_ctx*, int dtype, float Krho, const sphb200_group_view* grp_h, const void* divv,
                                float* groupDt);
int sphb200_group_acc_timestep(sphb200_ctx*, int dtype, float etaAcc, const sphb200_group_view* grp_h, const void* ax,
                               const void* ay, const void* az, float* groupDt);
/* the group minimum momentumAndEnergyGpu folds into groupDt (hydro_ve/momentum_energy_gpu.cu:98-104), from the
 * per-particle Courant steps sphb200_momentum_energy_targets leaves in dtParticle (indexed like the fields) */
int sphb200_group_min_timestep(sphb200_ctx*, int dtype, const sphb200_group_view* grp_h, const void* dtParticle,
                               float* groupDt);
/* storeRungGpu (ts_groups.cu:80-103) */
int sphb200_store_rung(sphb200_ctx*, const sphb200_group_view* grp_h, uint8_t rung, uint8_t* particleRungs);
/* extractGroupGpu (sph/include/sph/groups.hpp:32-47): out[k] = grp[indices[first + k]], k < last - first */
int sphb200_extract_groups(sphb200_ctx*, const sphb200_group_view* grp_h, const uint32_t* indices, uint32_t first,
                           uint32_t last, uint32_t* outStart, uint32_t* outEnd);
/* the particles of a set of groups as an ascending index list: what the *_targets operators below take.
 * targets: device, room for lastBody - firstBody entries */
int sphb200_targets_from_groups(sphb200_ctx*, const sphb200_group_view* grp_h, uint32_t* targets, uint32_t* numTargets_h);

/* GroupView scope of the target-range operators.  The reference's GPU operators take a GroupView (sph_gpu.hpp:13-82)
 * and the tree view carries searchExtFactor (octree.hpp:315); here both are context state: while a scope is set,
 * sphb200_find_neighbors, _xmass, _ve_def_gradh, _iad_divv_curlv, _av_switches and _momentum_energy only process the
 * targets i of [first,last) with active[i] != 0 (indexed like the fields) and leave every output of the others
 * untouched.  NULL resets (all targets, factor 1, no per-target Courant output). */
typedef struct sphb200_op_scope
{
    const uint8_t* active;  /* device, NULL = all of [first,last) */
    float searchExtFactor;  /* <= 0: 1.  > 1 while particles have drifted since the tree was built */
    void* dtParticle;       /* device T[...] indexed like the fields or NULL: Courant step of every processed target,
                               written by sphb200_momentum_energy */
    uint8_t* gatherMask;    /* device, one byte per particle (f->n), zeroed by the caller, or NULL.  With `active` set,
                               sphb200_find_neighbors marks every target it processes and every particle on its list;
                               the pair operators then refresh their gather records for the marked particles only
                               (a partial sync touches a few per cent of them) */
} sphb200_op_scope;
int sphb200_set_scope(sphb200_ctx*, const sphb200_op_scope* scope_h);
/* active[0..n) = 0, then active[i] = 1 for the members of the groups */
int sphb200_mark_groups(sphb200_ctx*, const sphb200_group_view* grp_h, uint32_t n, uint8_t* active);

/* sortGroupDt (sph/include/sph/ts_rungs.hpp:55-66): stable ascending sort of groupDt[0..numGroups) in place,
 * groupIndices[k] = previous position of the k-th smallest */
int sphb200_sort_group_dt(sphb200_ctx*, float* groupDt, uint32_t* groupIndices, uint32_t numGroups);
/* cstone::sequenceGpu (primitives_gpu.h:55): out[k] = first + k */
int sphb200_sequence(sphb200_ctx*, uint32_t* out, uint32_t n, uint32_t first);
/* findRungRanges (ts_rungs.hpp:115-129) on sorted groupDt: rungRanges_h[0] = 0, [r] = lower_bound(2^r minDt) for
 * r < numRungs, numGroups for the rest (5 entries); minDt < 0: use groupDt[0].  minGroupDt_h (may be NULL) =
 * timestepRangeGpu (:69-75): {groupDt[0], groupDt[uint32(fastFraction * numGroups)]} */
int sphb200_rung_ranges(sphb200_ctx*, float minDt, const float* groupDt, uint32_t numGroups, int numRungs,
                        float fastFraction, uint32_t* rungRanges_h, float* minGroupDt_h);

/* driftPositionsGpu (sph/include/sph/positions_gpu.cu:47-111): drift the groups' particles back by dt_back to where
 * their accelerations were evaluated, then forward by dt: x,y,z,v,temp rewritten, x_m1/du_m1 kept, no fold into the
 * box.  dt_m1_h: 4 floats (Timestep::dt_m1), rung: device per-particle rung or NULL (= rung 0) */
int sphb200_drift_positions(sphb200_ctx*, int dtype, const sphb200_group_view* grp_h, float dt, float dt_back,
                            const float* dt_m1_h, const uint8_t* rung, const sphb200_hydro_const* hc_h,
                            const sphb200_fields* f_h);
/* computePositionsGpu (positions_gpu.cu:118-180): advance the groups' particles by dt (x,y,z,v,x_m1,temp,du_m1) */
int sphb200_compute_positions_groups(sphb200_ctx*, int dtype, const sphb200_group_view* grp_h, float dt,
                                     const float* dt_m1_h, const uint8_t* rung, const sphb200_box* box_h,
                                     const sphb200_hydro_const* hc_h, const sphb200_fields* f_h);
/* updateSmoothingLengthGpu (sph/include/sph/update_h_gpu.cu:39-60); nc indexed like h */
int sphb200_update_h_groups(sphb200_ctx*, int dtype, const sphb200_group_view* grp_h, unsigned ng0, const uint32_t* nc,
                            void* h);

/* ---- HydroVeBdtProp for one rank on a sphb200_sim (main/src/propagator/ve_hydro_bdt.hpp:51-420, g = 0) ----
 * sphb200_sim_bdt_enable switches the sim to block time steps: per-group time steps sorted into <= 4 power-of-two
 * rungs; a hierarchy of 2^(numRungs-1) substeps of length nextDt starts with a full sync (domain sync + new spatial
 * groups), the substeps in between only touch the groups of the rungs that are due (partialSync: no re-sort, search
 * extent factor x1.012 per substep), every other particle is drifted.  safetySteps > 0 forces numRungs = 1 for that
 * many hierarchies (HydroVeBdtProp::load after a split restart). */
typedef struct sphb200_bdt_info
{
    float nextDt, elapsedDt, totDt;
    int numRungs, substep;
    uint32_t rungRanges[5];
    float dt_m1[4], dt_drift[4];
    uint32_t numGroups;       /* spatial groups of the last full sync */
    uint32_t numActiveGroups; /* groups whose forces were evaluated in the last computeForces */
    float searchExtFactor;
    int isSynced; /* Propagator::isSynced: the next computeForces starts with a full sync */
} sphb200_bdt_info;
/* etaAcc, eps <= 0: ParticlesData defaults 0.2 and 0.005 (acceleration criterion etaAcc * sqrt(eps) / sqrt(|a|)) */
int sphb200_sim_bdt_enable(sphb200_sim*, int safetySteps, double etaAcc, double eps);
int sphb200_sim_bdt_compute_forces(sphb200_sim*); /* sync (full or partial) + forces of the active rungs */
int sphb200_sim_bdt_integrate(sphb200_sim*);      /* computeRungs + advance/drift + updateSmoothingLength */
int sphb200_sim_bdt_step(sphb200_sim*);
int sphb200_sim_bdt_info(const sphb200_sim*, sphb200_bdt_info* out);
/* restart support: the time-step record a new hierarchy starts from (substep 0), and the rung of the particles of
 * the last sphb200_sim_upload [0,n) / of the owned range */
int sphb200_sim_bdt_restore(sphb200_sim*, int numRungs, const float dt_m1[4]);
int sphb200_sim_bdt_upload_rungs(sphb200_sim*, const uint8_t* host, size_t n);
int sphb200_sim_bdt_download_rungs(sphb200_sim*, uint8_t* host);
/* device views for tests: per-particle rung [nWithHalos], groupDt [numGroups], and the active-target mask */
const uint8_t* sphb200_sim_bdt_rungs(const sphb200_sim*);
const float* sphb200_sim_bdt_group_dt(const sphb200_sim*);
const uint8_t* sphb200_sim_bdt_active(const sphb200_sim*);
/* groups of the last full sync in time-step order (tsGroups_): start/end device arrays of info.numGroups entries */
int sphb200_sim_bdt_groups(const sphb200_sim*, sphb200_group_view* out);

#ifdef __cplusplus
}
#endif
#endif

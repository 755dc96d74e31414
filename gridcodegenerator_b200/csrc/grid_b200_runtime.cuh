// grid_b200_runtime.cuh -- hand-written sm_100a runtime shared by every emitted grid.cuh.
//
// Execution model (replaces the reference's "one thread block per knot point, ~40 __syncthreads per
// knot", helpers/_code_generation_helpers.py:41-67, algorithms/_forward_dynamics_gradient.py:113-184):
//   * one lane owns one knot point, one warp owns a tile of 32 knot points; warps exchange no data -- the
//     only CTA barriers are lockstep points that keep the warps of an SM on the same instruction-cache lines;
//   * the tile's inputs (32 x stride contiguous values of the reference's AoS batch, e.g. [q|qd|u]) are
//     fetched by ONE TMA bulk copy (cp.async.bulk + mbarrier), double buffered so the next tile's load
//     overlaps this tile's math; the tile's outputs are assembled in shared memory and written by ONE
//     TMA bulk store -> HBM traffic is exactly the algorithmic bytes, fully coalesced;
//   * per-thread pass-to-pass state lives in shared memory as S[slot*NT + tid] (conflict-free,
//     immediate-offset LDS/STS); robots whose state does not fit use a global scratch with the same layout.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include <mutex>
#ifndef GRID_PDL_PREFETCH
#define GRID_PDL_PREFETCH 1
#endif
#ifndef GRID_PDL_DEFAULT
#define GRID_PDL_DEFAULT 1
#endif

// compiler-only memory fence used by the emitted bodies at joint / column boundaries
#define GRID_FENCE() asm volatile("" ::"l"(S) : "memory")
// An opaque copy of a scalar: the second-order bodies run the RNEA once per seed-pair block on laundered copies of
// their inputs, so the compiler cannot merge the value-level work of different passes (and keep it live across the
// whole kernel).  Costs at most a register move.
__device__ __forceinline__ float grid_launder(float x) { asm volatile("" : "+f"(x)); return x; }
__device__ __forceinline__ double grid_launder(double x) { asm volatile("" : "+d"(x)); return x; }
#define GRID_LAUNDER(x) grid_launder(x)
// Emitted once per body before its first output write (see tile_loop: the output staging buffer of a TMA-staged
// tile is still being read by the previous tile's bulk store when the next body starts; the body's first passes
// write no output, so the wait is deferred to this point and normally finds the store long finished).
#ifndef GRID_BEFORE_FIRST_OUTPUT
#define GRID_BEFORE_FIRST_OUTPUT() grid_rt::output_buffer_acquire()
#endif

namespace grid_rt {

// Accurate sincos (the reference emits double-precision sin/cos then casts, helpers/_topology_helpers.py:194-195; the
// fast __sincosf would not hold 1e-4 through a depth-7 chain).  By default ONE out-of-line copy per kernel:
// inlined, every joint carries its own range-reduction slow path (~100 SASS instructions each), which pushes
// that much of the straight-line body beyond the 64 KB that the instruction cache serves without stalls.
#ifndef GRID_SINCOS_INLINE
static __device__ __noinline__ float2 grid_sincos_call(float x) { float s, c; sincosf(x, &s, &c); return make_float2(s, c); }
static __device__ __noinline__ double2 grid_sincos_call(double x) { double s, c; sincos(x, &s, &c); return make_double2(s, c); }
__device__ __forceinline__ void grid_sincos(float x, float* s, float* c) { const float2 r = grid_sincos_call(x); *s = r.x; *c = r.y; }
__device__ __forceinline__ void grid_sincos(double x, double* s, double* c) { const double2 r = grid_sincos_call(x); *s = r.x; *c = r.y; }
#else
__device__ __forceinline__ void grid_sincos(float x, float* s, float* c) { sincosf(x, s, c); }
__device__ __forceinline__ void grid_sincos(double x, double* s, double* c) { sincos(x, s, c); }
#endif

// ---- pairs of scalars: the dq / dqd derivative columns of a joint go through identical linear maps, so the
// emitted bodies carry them as one P value.  For T = float P is float2 and every operation below is ONE packed
// instruction on sm_100a (FFMA2 / FMUL2 / FADD2 with the scalar coefficient as a broadcast .F32 operand), which
// halves the instruction stream of the gradient passes; for T = double it is two scalar operations.
struct double_pair { double x, y; };
template <typename T> struct pair_of;
template <> struct pair_of<float> { typedef float2 type; };
template <> struct pair_of<double> { typedef double_pair type; };
template <typename T> using pair_t = typename pair_of<T>::type;

template <typename T> __device__ __forceinline__ pair_t<T> mkp(T a, T b);
template <> __device__ __forceinline__ float2 mkp<float>(float a, float b) { return make_float2(a, b); }
template <> __device__ __forceinline__ double_pair mkp<double>(double a, double b) { double_pair r; r.x = a; r.y = b; return r; }
template <typename T> __device__ __forceinline__ pair_t<T> pbc(T s) { return mkp<T>(s, s); }

#ifdef GRID_PAIR_SCALAR   // A/B switch: the two halves as independent scalar FP32 operations (interleaved chains)
__device__ __forceinline__ float2 pfma(float2 a, float s, float2 c) { return make_float2(fmaf(a.x, s, c.x), fmaf(a.y, s, c.y)); }
__device__ __forceinline__ float2 pmul(float2 a, float s) { return make_float2(a.x * s, a.y * s); }
__device__ __forceinline__ float2 padd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 psub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ float2 padds(float2 a, float s) { return make_float2(a.x + s, a.y + s); }
#else
__device__ __forceinline__ float2 pfma(float2 a, float s, float2 c) { return __ffma2_rn(a, make_float2(s, s), c); }
__device__ __forceinline__ float2 pmul(float2 a, float s) { return __fmul2_rn(a, make_float2(s, s)); }
__device__ __forceinline__ float2 padd(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ float2 psub(float2 a, float2 b) { return __ffma2_rn(b, make_float2(-1.0f, -1.0f), a); }
__device__ __forceinline__ float2 padds(float2 a, float s) { return __fadd2_rn(a, make_float2(s, s)); }
#endif
__device__ __forceinline__ float2 pneg(float2 a) { return make_float2(-a.x, -a.y); }

__device__ __forceinline__ double_pair pfma(double_pair a, double s, double_pair c) { double_pair r; r.x = fma(a.x, s, c.x); r.y = fma(a.y, s, c.y); return r; }
__device__ __forceinline__ double_pair pmul(double_pair a, double s) { double_pair r; r.x = a.x * s; r.y = a.y * s; return r; }
__device__ __forceinline__ double_pair padd(double_pair a, double_pair b) { double_pair r; r.x = a.x + b.x; r.y = a.y + b.y; return r; }
__device__ __forceinline__ double_pair psub(double_pair a, double_pair b) { double_pair r; r.x = a.x - b.x; r.y = a.y - b.y; return r; }
__device__ __forceinline__ double_pair padds(double_pair a, double s) { double_pair r; r.x = a.x + s; r.y = a.y + s; return r; }
__device__ __forceinline__ double_pair pneg(double_pair a) { double_pair r; r.x = -a.x; r.y = -a.y; return r; }

constexpr int TILE = 32;          // knot points per warp tile
constexpr int MAX_STREAMS = 3;    // input arrays per algorithm (e.g. q_qd_u, qdd, Minv)

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// TMA 1-D bulk copies (no tensor map needed: a tile of the AoS batch is one contiguous byte range)
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* dst_gmem, const void* src_smem, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                 ::"l"(dst_gmem), "r"(smem_u32(src_smem)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// Programmatic dependent launch (PDL).  The kernels here are short (5-30 us), so the launch latency and the
// prologue (shared-memory init, mbarrier init) of kernel i+1 are worth hiding behind kernel i.  Every tile kernel
// releases its dependents at once (pdl_release) and blocks at pdl_wait before its first global-memory access;
// griddepcontrol.wait returns only when every prerequisite grid has completed and its writes are visible, so
// stream order semantics are unchanged.  Both are no-ops when the kernel was launched without the attribute
// (GRID_PDL=0 in the environment, or a plain <<<>>> launch of X_kernel by user code).
__device__ __forceinline__ void pdl_release() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// lane 0 issued the previous bulk store of this warp's staging buffer; no-op when no bulk group is outstanding
__device__ __forceinline__ void output_buffer_acquire() {
    if ((threadIdx.x & 31) == 0) bulk_wait_read0();
    __syncwarp();
}

template <typename T>
struct TileArgs {
    const T* in[MAX_STREAMS];   // device arrays, knot-major
    int stride[MAX_STREAMS];    // elements between consecutive knot points
    T* out;                     // OUT elements per knot point, contiguous
    T* scratch;                 // global state scratch (only when the state is not in shared memory)
    T gravity;
    int num_knots;
};

__host__ __device__ constexpr size_t align16(size_t x) { return (x + 15) & ~size_t(15); }
// per-knot element count of input stream s of algorithm A (0 = stream unused)
template <class A>
__host__ __device__ constexpr int in_n(int s) { return s == 0 ? A::IN0 : (s == 1 ? A::IN1 : A::IN2); }

// Shared-memory footprint (bytes) of one CTA of NT threads.
template <class A, typename T, int NT, bool STATE_SMEM, int STAGED>
constexpr size_t tile_smem_bytes() {
    size_t state = STATE_SMEM ? align16(size_t(A::SLOTS) * NT * sizeof(T)) : 0;
    size_t per_warp = 0;
    if (STAGED == 3) return state + size_t(TILE) * 33 * sizeof(T) * (NT / 32) + 16;   // one transpose tile per warp
    if (STAGED) {
        for (int s = 0; s < MAX_STREAMS; s++) per_warp += STAGED * align16(size_t(TILE) * in_n<A>(s) * sizeof(T));
        per_warp += align16(size_t(TILE) * A::OUT * sizeof(T));
        per_warp += 16;  // two mbarriers
    }
    return state + per_warp * (NT / 32) + 16;
}

// Lockstep point inside the emitted bodies: CTAs with several warps re-converge at every section boundary so
// that all warps of an SM stream the same straight-line code (>100 KB for a 7-DoF grad FD, far beyond the
// 32 KB L1.5 / 6 KB L0 instruction caches) through the instruction caches together: one fetch per SM instead of
// one per warp.  Measured on B200 (profiles/): with independent single-warp CTAs the kernel was
// instruction-fetch bound (stall "no_instruction" 3.1 cycles per issued instruction, issue slots 21 % busy).
// GRID_LOCKSTEP (set by the emitter; 0 when the bodies contain no barriers): with barriers every warp of the CTA
// must run every round, tile or not; without them a warp leaves as soon as it has no tile left, which frees issue
// slots in the last, partial round.
#ifndef GRID_LOCKSTEP
#define GRID_LOCKSTEP 1
#endif
// Only the warps that own a tile in the current round run the body; they meet at named barrier 1 with a thread
// count that tile_loop publishes per round (warps without a tile skip the round entirely: no dummy arithmetic,
// stores or spills -- this matters for small batches, where most warps of the persistent grid have no tile).
__device__ __forceinline__ int& sync_threads_ref() { __shared__ int n; return n; }
template <int NT>
__device__ __forceinline__ void cta_sync() {
    if (NT > 32) {
        const int n = *reinterpret_cast<volatile int*>(&sync_threads_ref());
        asm volatile("bar.sync 1, %0;" ::"r"(n));
    }
}
// called by every warp of the CTA at the top of a round; returns whether this warp has a tile
template <int NT>
__device__ __forceinline__ bool round_begin(int tile_of_warp0, int stride_w, int num_tiles, int warp) {
    if (NT > 32 && GRID_LOCKSTEP) {
        // tiles of warps 0..NWARPS-1 in this round are tile_of_warp0 + w*stride_w (increasing): count the valid ones
        int nact = 0;
        if (tile_of_warp0 < num_tiles) nact = min(NT / 32, (num_tiles - 1 - tile_of_warp0) / stride_w + 1);
        __syncthreads();                        // previous round's barrier-1 users are done with the old count
        if (threadIdx.x == 0) sync_threads_ref() = nact * 32;
        __syncthreads();
        return warp < nact;
    }
    return tile_of_warp0 + warp * stride_w < num_tiles;
}
#define GRID_SYNC() grid_rt::cta_sync<NT>()

// A CTA of NT threads (NT/32 warps) walks the batch in rounds; in each round warp w of CTA b owns tile
// (round*(NT/32) + w)*gridDim.x + b (32 knot points, one per lane).  Warps without a tile in a round skip it; the
// lockstep barriers inside the body count only the warps that run (round_begin / cta_sync).
// The kernel must be launched with exactly NT threads per block.
// STAGED: 0 = plain loads / stores; 2 = TMA-staged with double-buffered inputs (next tile's load overlaps this
// tile's math); 1 = TMA-staged with one input buffer (next load issued after the body; frees shared memory for
// more lockstep warps when the batch fits one or two rounds anyway); 3 = transposed store for large per-knot
// outputs: plain input loads, each lane writes its output row into a lane-interleaved global scratch (element i of
// lane l at [i*32 + l], one full 128-byte line per store instruction), then the warp copies the scratch into the
// caller's knot-major rows through a 32 x 33 shared-memory transpose tile (coalesced on both sides).
template <class A, typename T, int NT, bool STATE_SMEM, int STAGED>
__device__ __forceinline__ void tile_loop(const TileArgs<T>& a, unsigned char* smem) {
    constexpr int NWARPS = NT / 32;
    if (blockDim.x != NT) return;                  // launch contract: exactly NT threads per CTA
    const int tid = threadIdx.x;
    const int lane = tid & 31, warp = tid >> 5;
    const int num_tiles = (a.num_knots + TILE - 1) / TILE;
    const int tiles_per_round = gridDim.x * NWARPS;
    const int rounds = (num_tiles + tiles_per_round - 1) / tiles_per_round;
    // warp w of CTA b owns tiles (r*NWARPS + w)*gridDim.x + b: the tiles of the last, partial round spread evenly
    // over the SMs (every SM keeps ~the same number of busy warps) instead of leaving whole SMs idle
    const int first = warp * gridDim.x + blockIdx.x;
    pdl_release();

    T* S;
    size_t off = 0;
    if (STATE_SMEM) {
        S = reinterpret_cast<T*>(smem) + tid;
        off = align16(size_t(A::SLOTS) * NT * sizeof(T));
    } else {
        S = a.scratch + size_t(blockIdx.x) * (size_t(NT) * A::SLOTS + (STAGED == 3 ? size_t(NT) * A::OUT : 0)) + tid;
    }

    if (STAGED == 0 || STAGED == 3) pdl_wait();     // the staged path waits after its shared-memory prologue
    if (STAGED == 3) {
        T* tile_s = reinterpret_cast<T*>(smem + off) + size_t(warp) * TILE * 33;
        T* scr = a.scratch + size_t(blockIdx.x) * ((STATE_SMEM ? 0 : size_t(NT) * A::SLOTS) + size_t(NT) * A::OUT)
                 + (STATE_SMEM ? 0 : size_t(NT) * A::SLOTS) + size_t(warp) * TILE * A::OUT;
        constexpr int NCHUNK = (A::OUT + TILE - 1) / TILE;
        T* scr0 = scr - size_t(warp) * TILE * A::OUT;          // warp 0's output scratch
        for (int r = 0; r < rounds; r++) {
            const int tile0 = blockIdx.x + r * tiles_per_round;  // warp 0's tile; warp w owns tile0 + w*gridDim.x
            const int nact = tile0 < num_tiles ? min(NWARPS, (num_tiles - 1 - tile0) / int(gridDim.x) + 1) : 0;
            if (round_begin<NT>(tile0, gridDim.x, num_tiles, warp)) {
                int k = (first + r * tiles_per_round) * TILE + lane;
                if (k >= a.num_knots) k = a.num_knots - 1;   // dummy lanes redo the last knot (identical values)
                const T* x[MAX_STREAMS];
#pragma unroll
                for (int s = 0; s < MAX_STREAMS; s++)
                    x[s] = in_n<A>(s) ? a.in[s] + size_t(k) * a.stride[s] : nullptr;
                A::template run<NT, TILE>(scr + lane, x[0], x[1], x[2], S, a.gravity);
            }
            if (NWARPS > 1) __syncthreads();                 // every active warp's scratch rows are written
            else __syncwarp();
            // All warps of the CTA (also those without a tile this round) share the transposition: work item p =
            // (source warp, chunk of 32 output elements); 32 independent scratch loads (L2 hits) in flight per lane.
            for (int p = warp; p < nact * NCHUNK; p += NWARPS) {
                const int ws = p / NCHUNK, i0 = (p - ws * NCHUNK) * TILE;
                const int tile = tile0 + ws * int(gridDim.x);
                const int nvalid = min(TILE, a.num_knots - tile * TILE);
                const T* src = scr0 + size_t(ws) * TILE * A::OUT + size_t(i0) * TILE + lane;
                T v[TILE];
#pragma unroll
                for (int i = 0; i < TILE; i++) v[i] = (i0 + i < A::OUT) ? __ldcg(src + i * TILE) : T(0);
#pragma unroll
                for (int i = 0; i < TILE; i++) tile_s[i * 33 + lane] = v[i];
                __syncwarp();
                if (i0 + lane < A::OUT) {
                    T* d = a.out + size_t(tile) * TILE * A::OUT + i0 + lane;
#pragma unroll 8
                    for (int l = 0; l < nvalid; l++) d[size_t(l) * A::OUT] = tile_s[lane * 33 + l];
                }
                __syncwarp();
            }
            if (NWARPS > 1) __syncthreads();                 // scratch rows are free for the next round
        }
        return;
    }

    if (!STAGED) {
        for (int r = 0; r < rounds; r++) {
            const int tile = first + r * tiles_per_round;
            if (!round_begin<NT>(blockIdx.x + r * tiles_per_round, gridDim.x, num_tiles, warp)) continue;
            int k = tile * TILE + lane;
            if (k >= a.num_knots) k = a.num_knots - 1;   // dummy lanes redo the last knot (identical values)
            const T* x[MAX_STREAMS];
#pragma unroll
            for (int s = 0; s < MAX_STREAMS; s++) x[s] = in_n<A>(s) ? a.in[s] + size_t(k) * a.stride[s] : nullptr;
            A::template run<NT, 1>(a.out + size_t(k) * A::OUT, x[0], x[1], x[2], S, a.gravity);
        }
        return;
    }

    // ---- staged path: carve this warp's buffers
    size_t per_warp = 0, in_off[MAX_STREAMS][2];
#pragma unroll
    for (int s = 0; s < MAX_STREAMS; s++)
#pragma unroll
        for (int b = 0; b < 2; b++) {
            in_off[s][b] = (b < STAGED) ? per_warp : in_off[s][0];
            if (b < STAGED) per_warp += align16(size_t(TILE) * in_n<A>(s) * sizeof(T));
        }
    const size_t out_off = per_warp;
    per_warp += align16(size_t(TILE) * A::OUT * sizeof(T));
    const size_t bar_off = per_warp;
    per_warp += 16;
    unsigned char* wbase = smem + off + per_warp * warp;
    T* out_s = reinterpret_cast<T*>(wbase + out_off);
    uint64_t* bar = reinterpret_cast<uint64_t*>(wbase + bar_off);

    // a tile can go through TMA when it is full and every stream is dense and 16-byte aligned
    bool dense = (reinterpret_cast<uintptr_t>(a.out) & 15) == 0;
#pragma unroll
    for (int s = 0; s < MAX_STREAMS; s++)
        if (in_n<A>(s)) dense = dense && a.stride[s] == in_n<A>(s) && (reinterpret_cast<uintptr_t>(a.in[s]) & 15) == 0;

    // While the previous kernel of the stream is still finishing (programmatic dependent launch), pull this warp's
    // first input tile into L2: a prefetch moves no data into the SM and L2 is the coherence point, so it is safe
    // even if that kernel is still writing the buffer; the TMA load after pdl_wait() then hits L2 instead of HBM.
    if (GRID_PDL_PREFETCH && dense && lane == 0 && first < num_tiles && (first + 1) * TILE <= a.num_knots) {
#pragma unroll
        for (int s = 0; s < MAX_STREAMS; s++)
            if (in_n<A>(s))
                asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(a.in[s] + size_t(first) * TILE * in_n<A>(s)),
                             "r"(uint32_t(TILE * in_n<A>(s) * sizeof(T))) : "memory");
    }
    // dummy lanes / dummy warps read whatever is in the staging buffers: make it finite
    for (size_t i = lane; i < out_off / sizeof(T); i += 32) reinterpret_cast<T*>(wbase)[i] = T(0);
    if (lane == 0) { mbar_init(&bar[0], 1); mbar_init(&bar[1], 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    fence_proxy_async();
    __syncwarp();
    pdl_wait();                                      // first global access (TMA load / plain load) is below

    auto issue_load = [&](int tile, int buf) {      // lane 0 only
        uint32_t bytes = 0;
#pragma unroll
        for (int s = 0; s < MAX_STREAMS; s++) bytes += in_n<A>(s) ? uint32_t(TILE * in_n<A>(s) * sizeof(T)) : 0u;
        mbar_expect_tx(&bar[buf], bytes);
#pragma unroll
        for (int s = 0; s < MAX_STREAMS; s++)
            if (in_n<A>(s))
                bulk_g2s(wbase + in_off[s][buf], a.in[s] + size_t(tile) * TILE * in_n<A>(s),
                         uint32_t(TILE * in_n<A>(s) * sizeof(T)), &bar[buf]);
    };
    auto tile_is_tma = [&](int tile) { return dense && tile < num_tiles && (tile + 1) * TILE <= a.num_knots; };

    uint32_t phase[2] = {0u, 0u};
    int buf = 0;
    if (tile_is_tma(first) && lane == 0) issue_load(first, 0);

    for (int r = 0; r < rounds; r++, buf = (STAGED == 2) ? (buf ^ 1) : 0) {
        const int tile = first + r * tiles_per_round;
        const int next = tile + tiles_per_round;
        const bool valid = tile < num_tiles;
        if (!round_begin<NT>(blockIdx.x + r * tiles_per_round, gridDim.x, num_tiles, warp)) continue;   // == !valid
        const bool tma = tile_is_tma(tile);
        if (STAGED == 2 && tile_is_tma(next) && lane == 0) issue_load(next, buf ^ 1);
        const int nvalid = valid ? min(TILE, a.num_knots - tile * TILE) : 0;
        const T* x[MAX_STREAMS];
#pragma unroll
        for (int s = 0; s < MAX_STREAMS; s++) x[s] = reinterpret_cast<const T*>(wbase + in_off[s][buf]) + lane * in_n<A>(s);
        if (tma) {
            mbar_wait(&bar[buf], phase[buf]);
            phase[buf] ^= 1u;
        } else if (valid) {                           // ragged tail / strided or unaligned caller buffers
#pragma unroll
            for (int s = 0; s < MAX_STREAMS; s++) {
                if (!in_n<A>(s)) continue;
                T* dst = reinterpret_cast<T*>(wbase + in_off[s][buf]);
                for (int i = lane; i < nvalid * in_n<A>(s); i += 32) {
                    const int kk = i / in_n<A>(s), e = i - kk * in_n<A>(s);
                    dst[i] = a.in[s][(size_t(tile) * TILE + kk) * a.stride[s] + e];
                }
            }
            __syncwarp();
        }
        // out_s may still be read by the previous tile's bulk store: the body waits for it right before its first
        // output write (GRID_BEFORE_FIRST_OUTPUT -> output_buffer_acquire), not here
        A::template run<NT, 1>(out_s + lane * A::OUT, x[0], x[1], x[2], S, a.gravity);
        __syncwarp();
        if (STAGED == 1 && tile_is_tma(next)) {      // single input buffer: every lane is done reading it
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) issue_load(next, 0);
        }
        if (tma) {
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) {
                bulk_s2g(a.out + size_t(tile) * TILE * A::OUT, out_s, uint32_t(TILE * A::OUT * sizeof(T)));
                bulk_commit();
            }
        } else if (valid) {
            T* dst = a.out + size_t(tile) * TILE * A::OUT;
            for (int i = lane; i < nvalid * A::OUT; i += 32) dst[i] = out_s[i];
            __syncwarp();
        }
    }
    if (lane == 0) bulk_wait_read0();                // smem must outlive the last bulk store's reads
    __syncwarp();
}

// ------------------------------------------------------------------ split (warp-specialised) tile loop
// One CTA of A::NW warps owns ONE 32-knot tile at a time (lane = knot in every warp).  Phase A (the first-order
// passes) runs on A::NA warps -- part p on warp p: the parts are independent pieces of the first-order work, e.g. the
// base-rooted trees (limbs) of the robot -- and publishes the per-knot state in shared memory as S[slot*32 + lane];
// after a CTA barrier every warp runs its share of the derivative columns (phase B: A::phaseB switches on the warp
// index) reading that state, writing the tile's output rows.  Against the one-lane-one-knot loop above this cuts the
// latency of a tile from the whole body to (longest phase-A part + longest column group) and spreads a tile's
// instruction stream over several instruction-fetch streams, which is what small batches (one tile per SM or
// fewer) and big robots (Atlas: 870 KB of straight-line code per tile) need.
// Shared-memory layout: [state A::SLOTS x 32][inputs, double buffered][output rows 32 x OUT if OUT_STAGED][2 mbarriers].
template <class A, typename T, bool OUT_STAGED>
constexpr size_t split_smem_bytes() {
    size_t b = align16(size_t(A::SLOTS) * TILE * sizeof(T));
    for (int s = 0; s < MAX_STREAMS; s++) b += A::NBUF * align16(size_t(TILE) * in_n<A>(s) * sizeof(T));
    if (OUT_STAGED) b += align16(size_t(TILE) * A::OUT * sizeof(T));
    return b + 32;
}

template <class A, typename T, bool OUT_STAGED>
__device__ __forceinline__ void tile_loop_split(const TileArgs<T>& a, unsigned char* smem) {
    constexpr int NW = A::NW;
    constexpr int NT = NW * 32;
    if (blockDim.x != NT) return;                  // launch contract: exactly NW warps per CTA
    const int tid = threadIdx.x;
    const int lane = tid & 31, warp = tid >> 5;
    const int num_tiles = (a.num_knots + TILE - 1) / TILE;
    pdl_release();

    T* S = reinterpret_cast<T*>(smem) + lane;
    size_t off = align16(size_t(A::SLOTS) * TILE * sizeof(T));
    size_t in_off[MAX_STREAMS][2];
#pragma unroll
    for (int s = 0; s < MAX_STREAMS; s++)
#pragma unroll
        for (int b = 0; b < 2; b++) {
            in_off[s][b] = (b < A::NBUF) ? off : in_off[s][0];
            if (b < A::NBUF) off += align16(size_t(TILE) * in_n<A>(s) * sizeof(T));
        }
    const size_t stage_end = off;
    T* out_s = reinterpret_cast<T*>(smem + off);
    if (OUT_STAGED) off += align16(size_t(TILE) * A::OUT * sizeof(T));
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + off);

    bool dense = (reinterpret_cast<uintptr_t>(a.out) & 15) == 0;
#pragma unroll
    for (int s = 0; s < MAX_STREAMS; s++)
        if (in_n<A>(s)) dense = dense && a.stride[s] == in_n<A>(s) && (reinterpret_cast<uintptr_t>(a.in[s]) & 15) == 0;

    const int first = blockIdx.x;
    if (GRID_PDL_PREFETCH && dense && tid == 0 && first < num_tiles && (first + 1) * TILE <= a.num_knots) {
#pragma unroll
        for (int s = 0; s < MAX_STREAMS; s++)
            if (in_n<A>(s))
                asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(a.in[s] + size_t(first) * TILE * in_n<A>(s)),
                             "r"(uint32_t(TILE * in_n<A>(s) * sizeof(T))) : "memory");
    }
    // dummy lanes of a ragged tile read whatever is in the input staging buffers: make it finite
    {
        T* z = reinterpret_cast<T*>(smem + in_off[0][0]);
        for (size_t i = tid; i < (stage_end - in_off[0][0]) / sizeof(T); i += NT) z[i] = T(0);
    }
    if (tid == 0) { mbar_init(&bar[0], 1); mbar_init(&bar[1], 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    fence_proxy_async();
    __syncthreads();
    pdl_wait();                                      // first global access is below

    auto issue_load = [&](int tile, int buf) {      // thread 0 only
        uint32_t bytes = 0;
#pragma unroll
        for (int s = 0; s < MAX_STREAMS; s++) bytes += in_n<A>(s) ? uint32_t(TILE * in_n<A>(s) * sizeof(T)) : 0u;
        mbar_expect_tx(&bar[buf], bytes);
#pragma unroll
        for (int s = 0; s < MAX_STREAMS; s++)
            if (in_n<A>(s))
                bulk_g2s(smem + in_off[s][buf], a.in[s] + size_t(tile) * TILE * in_n<A>(s),
                         uint32_t(TILE * in_n<A>(s) * sizeof(T)), &bar[buf]);
    };
    auto tile_is_tma = [&](int tile) { return dense && tile < num_tiles && (tile + 1) * TILE <= a.num_knots; };

    uint32_t phase[2] = {0u, 0u};
    int buf = 0;
    if (tid == 0 && tile_is_tma(first)) issue_load(first, 0);

    // A::NBUF input buffers: 2 = the next tile's load is issued at the top of this tile (overlaps phase A and B);
    // 1 = it is issued once phase A has consumed this tile's inputs (overlaps phase B only) -- chosen by the emitter
    // when the second buffer would cost a resident CTA per SM.
    for (int tile = first; tile < num_tiles; tile += gridDim.x, buf = (A::NBUF == 2) ? (buf ^ 1) : 0) {
        const int next = tile + gridDim.x;
        const bool tma = tile_is_tma(tile);
        const int nvalid = min(TILE, a.num_knots - tile * TILE);
        if (A::NBUF == 2 && tid == 0 && tile_is_tma(next)) issue_load(next, buf ^ 1);   // buf^1 was last read two barriers ago
        if (A::PHASE_A_WRITES && OUT_STAGED) {                // phase A writes output rows too (side outputs)
            if (tid == 0) bulk_wait_read0();
            __syncthreads();
        }
        int k = tile * TILE + lane;
        if (k >= a.num_knots) k = a.num_knots - 1;           // dummy lanes of a ragged tile redo its last knot
        T* o = OUT_STAGED ? out_s + size_t(lane) * A::OUT : a.out + size_t(k) * A::OUT;
        if (warp < A::NA) {
            const T* x[MAX_STREAMS];
            if (tma) {
                mbar_wait(&bar[buf], phase[buf]);
#pragma unroll
                for (int s = 0; s < MAX_STREAMS; s++)
                    x[s] = reinterpret_cast<const T*>(smem + in_off[s][buf]) + lane * in_n<A>(s);
            } else {                                          // ragged tail / strided or unaligned caller buffers
#pragma unroll
                for (int s = 0; s < MAX_STREAMS; s++)
                    x[s] = in_n<A>(s) ? a.in[s] + size_t(k) * a.stride[s] : nullptr;
            }
            A::template phaseA<TILE, 1>(warp, o, x[0], x[1], x[2], S, a.gravity);
            if (A::NBUF == 1) fence_proxy_async();            // generic reads of the input buffer before its next TMA write
        }
        if (tma) phase[buf] ^= 1u;
        // the output rows may still be read by the previous tile's bulk store (issued by thread 0)
        if (OUT_STAGED && tid == 0) bulk_wait_read0();
        __syncthreads();                                      // state published, output rows free
        if (A::NBUF == 1 && tid == 0 && tile_is_tma(next)) issue_load(next, 0);   // every phase-A warp is done with buffer 0
        A::template phaseB<TILE, 1>(warp, o, S);
        if (OUT_STAGED) {
            fence_proxy_async();
            __syncthreads();                                  // all columns written; state free for the next tile
            if (tma) {
                if (tid == 0) {
                    bulk_s2g(a.out + size_t(tile) * TILE * A::OUT, out_s, uint32_t(TILE * A::OUT * sizeof(T)));
                    bulk_commit();
                }
            } else {
                T* dst = a.out + size_t(tile) * TILE * A::OUT;
                for (int i = tid; i < nvalid * A::OUT; i += NT) dst[i] = out_s[i];
                __syncthreads();                              // out_s is rewritten by the next tile's phase A
            }
        } else {
            __syncthreads();
        }
    }
    if (OUT_STAGED && tid == 0) bulk_wait_read0();   // smem must outlive the last bulk store's reads
    __syncthreads();
}

// ------------------------------------------------------------------ split tile loop, several tiles per CTA pass
// Big robots (Atlas: 830 KB of straight-line SASS per tile) are bound by instruction supply: code beyond the ~128 KB
// the SM's instruction caches hold streams from L2 at ~2.7 B/clk per SM, however many warps run it
// (tools/probes/icache_stream.py, profiles/r02j_icache_stream.log), and in the split form every warp of a CTA runs
// DIFFERENT code, so a tile costs one pass over the whole body.  Here a CTA holds the published state of A::NTILE
// tiles in shared memory at once: phase A runs (tile, part) units concurrently on the CTA's warps (the same part of
// two tiles on two warps at the same time -> one fetch), and in phase B every column function loops over the
// resident tiles before the warp moves on to its next column, so a fetched column body (10-30 KB) is executed
// NTILE times out of the instruction cache.
// Output rows of such robots do not fit shared memory (Atlas: 225 KB per tile) and a lane writing its own row costs
// 32 sectors per store instruction, so every lane writes its row into a lane-interleaved global scratch (element i
// of lane l of tile t at [(t*OUT + i)*32 + l]: one full 128-byte line per store instruction, L2-resident) and after
// phase B the CTA copies the scratch into the caller's knot-major rows through 32 x 33 transpose tiles that reuse the
// (then dead) state area -- the same transposed store as tile_loop's STAGED == 3.
// Batches with fewer tiles than CTAs run one tile per pass (latency first).
template <class A, typename T>
constexpr size_t split_multi_smem_bytes() {
    size_t b = align16(size_t(A::NTILE) * A::SLOTS * TILE * sizeof(T));
    for (int s = 0; s < MAX_STREAMS; s++) b += align16(size_t(A::NTILE) * TILE * in_n<A>(s) * sizeof(T));
    return b + 32;
}
// scratch elements one CTA of tile_loop_split_multi addresses
template <class A>
constexpr size_t split_multi_scratch_elems() { return A::XPOSE ? size_t(A::NTILE) * TILE * A::OUT : 0; }

template <class A, typename T>
__device__ __forceinline__ void tile_loop_split_multi(const TileArgs<T>& a, unsigned char* smem) {
    constexpr int NW = A::NW;
    constexpr int NT = NW * 32;
    constexpr size_t STATE = size_t(A::SLOTS) * TILE;          // elements of one tile's published state
    constexpr size_t OROWS = size_t(TILE) * A::OUT;            // scratch elements of one tile's output rows
    constexpr int NCHUNK = (A::OUT + TILE - 1) / TILE;
    constexpr bool XP = A::XPOSE;                              // transposed store through the global scratch
    constexpr int OS = XP ? TILE : 1;                          // element stride of the rows the bodies write
    static_assert(A::NTILE <= 2 && NW % A::NTILE == 0, "phase B pairs the warps of a CTA over at most two tiles");
    static_assert(size_t(A::NTILE) * STATE >= size_t(NW) * TILE * 33, "transpose tiles reuse the state area");
    if (blockDim.x != NT) return;                  // launch contract: exactly NW warps per CTA
    const int tid = threadIdx.x;
    const int lane = tid & 31, warp = tid >> 5;
    const int num_tiles = (a.num_knots + TILE - 1) / TILE;
    const int G = num_tiles > int(gridDim.x) ? A::NTILE : 1;   // tiles per pass
    const int num_groups = (num_tiles + G - 1) / G;
    pdl_release();

    T* S = reinterpret_cast<T*>(smem) + lane;
    T* tile_s = reinterpret_cast<T*>(smem) + size_t(warp) * TILE * 33;
    T* scr = XP ? a.scratch + size_t(blockIdx.x) * split_multi_scratch_elems<A>() : nullptr;
    size_t off = align16(size_t(A::NTILE) * STATE * sizeof(T));
    size_t in_off[MAX_STREAMS];
#pragma unroll
    for (int s = 0; s < MAX_STREAMS; s++) {
        in_off[s] = off;
        off += align16(size_t(A::NTILE) * TILE * in_n<A>(s) * sizeof(T));
    }
    const size_t stage_end = off;
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + off);

    bool dense = true;
#pragma unroll
    for (int s = 0; s < MAX_STREAMS; s++)
        if (in_n<A>(s)) dense = dense && a.stride[s] == in_n<A>(s) && (reinterpret_cast<uintptr_t>(a.in[s]) & 15) == 0;

    {   // dummy lanes of a ragged group read whatever is in the input staging buffers: make it finite
        T* z = reinterpret_cast<T*>(smem + in_off[0]);
        for (size_t i = tid; i < (stage_end - in_off[0]) / sizeof(T); i += NT) z[i] = T(0);
    }
    if (tid == 0) mbar_init(&bar[0], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    fence_proxy_async();
    __syncthreads();
    pdl_wait();                                      // first global access is below

    auto group_is_tma = [&](int g) { return dense && g < num_groups && (g + 1) * G * TILE <= a.num_knots; };
    auto issue_load = [&](int g) {                  // thread 0 only: the G tiles of a group are contiguous in the batch
        uint32_t bytes = 0;
#pragma unroll
        for (int s = 0; s < MAX_STREAMS; s++) bytes += in_n<A>(s) ? uint32_t(G * TILE * in_n<A>(s) * sizeof(T)) : 0u;
        mbar_expect_tx(&bar[0], bytes);
#pragma unroll
        for (int s = 0; s < MAX_STREAMS; s++)
            if (in_n<A>(s))
                bulk_g2s(smem + in_off[s], a.in[s] + size_t(g) * G * TILE * in_n<A>(s),
                         uint32_t(G * TILE * in_n<A>(s) * sizeof(T)), &bar[0]);
    };

    uint32_t phase = 0u;
    const int first = blockIdx.x;
    if (tid == 0 && group_is_tma(first)) issue_load(first);

    for (int g = first; g < num_groups; g += gridDim.x) {
        const int next = g + gridDim.x;
        const bool tma = group_is_tma(g);
        const int nt = min(G, num_tiles - g * G);            // tiles present in this group
        // phase A: unit u = (tile t of the group, part p); the warps take units round-robin
        for (int u = warp; u < nt * A::NA; u += NW) {
            const int t = u / A::NA, part = u - t * A::NA;
            int k = (g * G + t) * TILE + lane;
            if (k >= a.num_knots) k = a.num_knots - 1;       // dummy lanes of a ragged tile redo the batch's last knot
            const T* x[MAX_STREAMS];
            if (tma) {
                mbar_wait(&bar[0], phase);
#pragma unroll
                for (int s = 0; s < MAX_STREAMS; s++)
                    x[s] = reinterpret_cast<const T*>(smem + in_off[s]) + size_t(t * TILE + lane) * in_n<A>(s);
            } else {                                          // ragged tail / strided or unaligned caller buffers
#pragma unroll
                for (int s = 0; s < MAX_STREAMS; s++)
                    x[s] = in_n<A>(s) ? a.in[s] + size_t(k) * a.stride[s] : nullptr;
            }
            A::template phaseA<TILE, OS>(part, XP ? scr + size_t(t) * OROWS + lane : a.out + size_t(k) * A::OUT, x[0], x[1], x[2],
                                         S + size_t(t) * STATE, a.gravity);
        }
        fence_proxy_async();                                  // generic reads of the input buffer before its next TMA write
        if (tma) phase ^= 1u;
        __syncthreads();                                      // state of every tile of the group published
        if (tid == 0 && group_is_tma(next)) issue_load(next);
        {
            // phase B: with two tiles resident, warps 2i and 2i+1 run the SAME columns on tile 0 and tile 1 at the same
            // time (one instruction fetch serves both); with one tile the two warps halve those columns between them
            int k0 = g * G * TILE + lane, k1 = k0 + TILE;
            if (k0 >= a.num_knots) k0 = a.num_knots - 1;
            if (k1 >= a.num_knots) k1 = a.num_knots - 1;
            T* o0 = XP ? scr + lane : a.out + size_t(k0) * A::OUT;
            T* o1 = XP ? scr + OROWS + lane : a.out + size_t(k1) * A::OUT;
            A::template phaseB_multi<TILE, OS>(warp, nt, o0, o1, S);
        }
        __syncthreads();                                      // every row of the group is written; state is dead
        // XPOSE: all warps share the transposition: work item p = (tile of the group, chunk of 32 output elements)
        for (int p = warp; XP && p < nt * NCHUNK; p += NW) {
            const int t = p / NCHUNK, i0 = (p - t * NCHUNK) * TILE;
            const int tile = g * G + t;
            const int nvalid = min(TILE, a.num_knots - tile * TILE);
            const T* src = scr + size_t(t) * OROWS + size_t(i0) * TILE + lane;
            T v[TILE];
#pragma unroll
            for (int i = 0; i < TILE; i++) v[i] = (i0 + i < A::OUT) ? __ldcg(src + i * TILE) : T(0);
#pragma unroll
            for (int i = 0; i < TILE; i++) tile_s[i * 33 + lane] = v[i];
            __syncwarp();
            if (i0 + lane < A::OUT) {
                T* d = a.out + size_t(tile) * TILE * A::OUT + i0 + lane;
#pragma unroll 8
                for (int l = 0; l < nvalid; l++) d[size_t(l) * A::OUT] = tile_s[lane * 33 + l];
            }
            __syncwarp();
        }
        __syncthreads();                                      // state area and scratch rows free for the next group
    }
}

// ------------------------------------------------------------------ host side
struct KernelPlan { const void* fn; int device; int blocks_per_sm; int sms; };

// Persistent grid for a tile kernel: resident CTAs per SM (from the occupancy calculator, after opting in
// to the large dynamic shared-memory carve-out) x SM count, clamped to the number of warp tiles.
// Cached per (kernel, device); returns <= 0 on error (cudaGetLastError() holds the reason).
// Internal linkage on purpose: several per-robot libraries live in one process (one .so per robot), and the
// local statics of an `inline` function would be merged across them by the dynamic linker (STB_GNU_UNIQUE).
static int persistent_blocks(const void* fn, int threads, size_t smem_bytes, int num_knots) {
    constexpr int MAX_PLANS = 256;
    static KernelPlan cache[MAX_PLANS];
    static int ncache = 0;
    static std::mutex mu;                    // the C ABI is callable from several host threads (one per device)
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return -1;
    std::lock_guard<std::mutex> lk(mu);
    const KernelPlan* plan = nullptr;
    for (int i = 0; i < ncache; i++)
        if (cache[i].fn == fn && cache[i].device == dev) { plan = &cache[i]; break; }
    if (!plan) {
        if (cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem_bytes)) != cudaSuccess) return -1;
        int per_sm = 0, sms = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, threads, smem_bytes) != cudaSuccess) return -1;
        if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return -1;
        if (per_sm <= 0) { per_sm = 0; }
        // GRID_MAX_BLOCKS_PER_SM (environment, tuning): cap on resident CTAs per SM.  Each CTA is one lockstep group,
        // i.e. one instruction stream; kernels whose straight-line body is far beyond the instruction caches are
        // bound by the SM's instruction-fetch path, which independent CTAs on one SM have to share.
        static const int env_cap = [] { const char* v = getenv("GRID_MAX_BLOCKS_PER_SM"); return (v && *v) ? atoi(v) : 0; }();
        if (env_cap > 0 && per_sm > env_cap) per_sm = env_cap;
        if (ncache >= MAX_PLANS) ncache = 0;         // recycle: a plan is cheap to recompute
        cache[ncache] = KernelPlan{fn, dev, per_sm, sms};
        plan = &cache[ncache++];
    }
    if (plan->blocks_per_sm <= 0) return 0;
    const int tiles = (num_knots + TILE - 1) / TILE;
    const int want = tiles;      // small batches: one tile per SM first (tile -> CTA mapping is round-robin over CTAs)
    (void)threads;
    const int cap = plan->blocks_per_sm * plan->sms;
    return want < cap ? want : cap;
}

// Launch with the programmatic-stream-serialization attribute unless GRID_PDL=0 (see pdl_release / pdl_wait).
static bool pdl_enabled() {
    static const int on = [] { const char* v = getenv("GRID_PDL"); return (v && *v) ? atoi(v) : GRID_PDL_DEFAULT; }();
    return on != 0;
}
template <typename... KArgs, typename... Args>
static cudaError_t launch_tile_kernel(void (*kern)(KArgs...), int blocks, int threads, size_t smem, cudaStream_t stream,
                                      Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(blocks); cfg.blockDim = dim3(threads); cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

// Which form of a kernel a launcher uses when the algorithm has both: GRID_SPLIT_MODE=0 never the split form,
// =1 always, otherwise the split form for batches of at most `max_knots` knot points (the latency-bound regime:
// too few tiles to give every SM's schedulers a warp; the emitted <algo>_split_max_knots() gives the default,
// GRID_SPLIT_MAX_KNOTS overrides it).
// The environment is read at every launch (two getenv calls, ~0.1 us) so that a process can switch forms, e.g.
// tests comparing the two.
static bool use_split_kernel(int num_knots, int max_knots) {
    const char* v = getenv("GRID_SPLIT_MODE");
    const int mode = (v && *v) ? atoi(v) : -1;
    v = getenv("GRID_SPLIT_MAX_KNOTS");
    const int env_max = (v && *v) ? atoi(v) : -1;
    if (mode == 0) return false;
    if (mode == 1) return true;
    return num_knots <= (env_max >= 0 ? env_max : max_knots);
}

// Number of CTAs the global state scratch attached to a robot model serves on the current device: a pure function
// of the device (SM count x blocks_per_sm), so that init_robotModel (which allocates it) and every launcher agree
// without sharing state across translation units.  Returns 0 on error.
static int scratch_capacity(int blocks_per_sm) {
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
    return sms * blocks_per_sm;
}

}  // namespace grid_rt

// mc.cu -- iso-surface extraction (marching cubes) over a dense SDF + mask grid.
//
// Reference: utils/mesher.py:310-337 `mc_mesh` = skimage.measure.marching_cubes(sdf, level=0, allow_degenerate=False, mask=...)
// on the CPU after copying the grid to the host, driven by recon_aabb_mesh / recon_aabb_collections_mesh (:396-485).
// scikit-image is an un-vendored dependency that is not installed here: the algorithm is the published one (Lorensen & Cline
// with a face-consistent resolution of ambiguous faces -- see oracle/mc_oracle.py, which restates the same rules on the CPU and
// is the parity partner of this file).  Grid layout = the mesher's: x-major, z fastest (utils/mesher.py:168-178).
//
//   k_mc_cells    one thread per grid point p = base corner of cell p: all 8 corners masked -> case index -> triangle count
//   k_mc_edges    one thread per grid point: which of its 3 owned edges (p -> p + e_axis) carries a vertex that an emitting cell uses
//   scans         exclusive prefix sums of the per-point vertex counts and per-cell triangle counts (3 small kernels each)
//   k_mc_emit     vertices (interpolated from the LOWER end point: bit-identical in every cell / rank that shares the edge) and
//                 indexed triangles, in grid order -> deterministic output, vertices de-duplicated by construction
// HBM-bound streaming kernels: 5 B read per point in the first pass (sdf + mask), 10 B/point of scratch.
#include <string.h>

#include <mutex>

#include "common.cuh"

namespace pinq {

constexpr int kMcMaxTris = 5;
__constant__ int8_t c_mc_tri[256][3 * kMcMaxTris + 1];  // edge ids, -1 padded
__constant__ uint8_t c_mc_ntri[256];

// ---- case table, derived on the host from the rules in the header of oracle/mc_oracle.py -------------------------------
// corner c = (c & 1, c >> 1 & 1, c >> 2 & 1); edge e = axis * 4 + (b1 + 2 b2), b1 / b2 = the other two coordinates
struct McTables {
    int8_t tri[256][3 * kMcMaxTris + 1];
    uint8_t ntri[256];
};

static inline void mc_edge_ends(int e, int& lo, int& hi) {
    const int a = e >> 2, k = e & 3;
    const int o0 = a == 0 ? 1 : 0, o1 = a == 2 ? 1 : 2;
    int bits[3] = {0, 0, 0};
    bits[o0] = k & 1;
    bits[o1] = k >> 1;
    lo = bits[0] | (bits[1] << 1) | (bits[2] << 2);
    bits[a] = 1;
    hi = bits[0] | (bits[1] << 1) | (bits[2] << 2);
}
static inline int mc_edge_between(int c0, int c1) {
    for (int e = 0; e < 12; ++e) {
        int lo, hi;
        mc_edge_ends(e, lo, hi);
        if ((lo == c0 && hi == c1) || (lo == c1 && hi == c0)) return e;
    }
    return -1;
}

static bool mc_build_tables(McTables& T) {
    int faces[6][4];
    int nf = 0;
    for (int a = 0; a < 3; ++a) {
        const int o0 = a == 0 ? 1 : 0, o1 = a == 2 ? 1 : 2;
        for (int side = 0; side < 2; ++side) {
            const int uv[4][2] = {{0, 0}, {1, 0}, {1, 1}, {0, 1}};
            for (int i = 0; i < 4; ++i) {
                int bits[3] = {0, 0, 0};
                bits[a] = side; bits[o0] = uv[i][0]; bits[o1] = uv[i][1];
                faces[nf][i] = bits[0] | (bits[1] << 1) | (bits[2] << 2);
            }
            ++nf;
        }
    }
    double mid[12][3];
    for (int e = 0; e < 12; ++e) {
        int lo, hi;
        mc_edge_ends(e, lo, hi);
        for (int d = 0; d < 3; ++d) mid[e][d] = 0.5 * (((lo >> d) & 1) + ((hi >> d) & 1));
    }
    for (int cs = 0; cs < 256; ++cs) {
        int nbr[12][2], deg[12];
        for (int e = 0; e < 12; ++e) deg[e] = 0;
        auto link = [&](int e0, int e1) {
            if (deg[e0] < 2) nbr[e0][deg[e0]] = e1;
            if (deg[e1] < 2) nbr[e1][deg[e1]] = e0;
            ++deg[e0]; ++deg[e1];
        };
        for (int f = 0; f < 6; ++f) {
            int cut[4], ncut = 0;
            for (int i = 0; i < 4; ++i) {
                const int c0 = faces[f][i], c1 = faces[f][(i + 1) & 3];
                if (((cs >> c0) & 1) != ((cs >> c1) & 1)) cut[ncut++] = mc_edge_between(c0, c1);
            }
            if (ncut == 2) link(cut[0], cut[1]);
            else if (ncut == 4) {  // ambiguous face: join the two cut edges that meet at each INSIDE corner
                for (int i = 0; i < 4; ++i)
                    if ((cs >> faces[f][i]) & 1)
                        link(mc_edge_between(faces[f][(i + 3) & 3], faces[f][i]), mc_edge_between(faces[f][i], faces[f][(i + 1) & 3]));
            }
        }
        for (int e = 0; e < 12; ++e)
            if (deg[e] != 0 && deg[e] != 2) return false;
        bool seen[12] = {false};
        int nt = 0;
        for (int e = 0; e <= 3 * kMcMaxTris; ++e) T.tri[cs][e] = -1;
        for (int start = 0; start < 12; ++start) {
            if (deg[start] == 0 || seen[start]) continue;
            int loop[12], n = 0, prev = -1, cur = start;
            loop[n++] = start;
            seen[start] = true;
            for (;;) {
                int nxt = (nbr[cur][0] != prev) ? nbr[cur][0] : nbr[cur][1];
                if (nbr[cur][0] == nbr[cur][1]) nxt = nbr[cur][0];
                if (nxt == start) break;
                loop[n++] = nxt;
                seen[nxt] = true;
                prev = cur;
                cur = nxt;
                if (n > 12) return false;
            }
            double nrm[3] = {0, 0, 0}, dir[3] = {0, 0, 0};
            for (int i = 0; i < n; ++i) {
                const double* p = mid[loop[i]];
                const double* q = mid[loop[(i + 1) % n]];
                nrm[0] += p[1] * q[2] - p[2] * q[1];
                nrm[1] += p[2] * q[0] - p[0] * q[2];
                nrm[2] += p[0] * q[1] - p[1] * q[0];
                int lo, hi;
                mc_edge_ends(loop[i], lo, hi);
                const int cin = ((cs >> lo) & 1) ? lo : hi, cout = ((cs >> lo) & 1) ? hi : lo;
                for (int d = 0; d < 3; ++d) dir[d] += ((cout >> d) & 1) - ((cin >> d) & 1);
            }
            const double s = nrm[0] * dir[0] + nrm[1] * dir[1] + nrm[2] * dir[2];
            if (s == 0.0) return false;
            if (s > 0)  // normals point towards decreasing values (skimage's default gradient_direction = "descent")
                for (int i = 0; i < n / 2; ++i) { const int t = loop[i]; loop[i] = loop[n - 1 - i]; loop[n - 1 - i] = t; }
            for (int i = 1; i + 1 < n; ++i) {
                if (nt >= kMcMaxTris) return false;
                T.tri[cs][3 * nt] = (int8_t)loop[0];
                T.tri[cs][3 * nt + 1] = (int8_t)loop[i];
                T.tri[cs][3 * nt + 2] = (int8_t)loop[i + 1];
                ++nt;
            }
        }
        T.ntri[cs] = (uint8_t)nt;
    }
    return true;
}

static const McTables* mc_tables() {
    static McTables T;
    static bool ok = false;
    static std::once_flag once;
    std::call_once(once, [] { ok = mc_build_tables(T); });
    return ok ? &T : nullptr;
}

static int mc_upload_tables(cudaStream_t st) {
    static bool uploaded[64] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && uploaded[dev]) return PINQ_OK;
    const McTables* T = mc_tables();
    if (!T) {
        set_error("marching-cubes case table could not be derived");
        return PINQ_ERR_CUDA;
    }
    cudaMemcpyToSymbolAsync(c_mc_tri, T->tri, sizeof(T->tri), 0, cudaMemcpyHostToDevice, st);
    cudaMemcpyToSymbolAsync(c_mc_ntri, T->ntri, sizeof(T->ntri), 0, cudaMemcpyHostToDevice, st);
    if (dev >= 0 && dev < 64) uploaded[dev] = true;
    return check_launch("marching-cubes table upload");
}

struct McGrid {
    int nx, ny, nz;
    int64_t n;
};

__device__ __forceinline__ int64_t mc_lin(const McGrid& g, int x, int y, int z) { return ((int64_t)x * g.ny + y) * g.nz + z; }

// cell p: bit 7 = all eight corners masked in (the cell may emit), low bits = triangle count
__global__ void k_mc_cells(const float* __restrict__ sdf, const uint8_t* __restrict__ mask, McGrid g, float level,
                           uint8_t* __restrict__ ctri) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= g.n) return;
    const int z = (int)(p % g.nz);
    const int64_t r = p / g.nz;
    const int y = (int)(r % g.ny), x = (int)(r / g.ny);
    uint8_t out = 0;
    if (x + 1 < g.nx && y + 1 < g.ny && z + 1 < g.nz) {
        bool ok = true;
        int cs = 0;
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            const int64_t q = mc_lin(g, x + (c & 1), y + ((c >> 1) & 1), z + ((c >> 2) & 1));
            ok = ok && (mask == nullptr || mask[q] != 0);
            cs |= (sdf[q] < level ? 1 : 0) << c;
        }
        if (ok) out = (uint8_t)(0x80u | c_mc_ntri[cs]);
    }
    ctri[p] = out;
}

// point p owns the edges p -> p + e_axis; bit `axis` of vflag = the edge carries a vertex (sign change) used by an emitting cell
__global__ void k_mc_edges(const float* __restrict__ sdf, const uint8_t* __restrict__ ctri, McGrid g, float level,
                           uint8_t* __restrict__ vflag) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= g.n) return;
    const int z = (int)(p % g.nz);
    const int64_t r = p / g.nz;
    const int y = (int)(r % g.ny), x = (int)(r / g.ny);
    const int c[3] = {x, y, z};
    const int dim[3] = {g.nx, g.ny, g.nz};
    const bool in0 = sdf[p] < level;
    uint8_t f = 0;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        if (c[a] + 1 >= dim[a]) continue;
        int q3[3] = {x, y, z};
        q3[a] += 1;
        if ((sdf[mc_lin(g, q3[0], q3[1], q3[2])] < level) == in0) continue;
        const int o0 = a == 0 ? 1 : 0, o1 = a == 2 ? 1 : 2;
        bool used = false;
#pragma unroll
        for (int k = 0; k < 4; ++k) {  // the four cells around the edge
            int b[3] = {x, y, z};
            b[o0] -= k & 1;
            b[o1] -= k >> 1;
            if (b[o0] < 0 || b[o1] < 0) continue;
            used = used || (ctri[mc_lin(g, b[0], b[1], b[2])] & 0x80u);
        }
        if (used) f |= (uint8_t)(1u << a);
    }
    vflag[p] = f;
}

// ---- exclusive scan of small per-element counts (uint8 in, int32 out); MODE 0: popcount of the 3 flag bits, 1: low 7 bits
constexpr int kScanThreads = 256, kScanPer = 8, kScanTileMc = kScanThreads * kScanPer;  // 2048 elements per block
template <int MODE>
__device__ __forceinline__ int mc_count_of(uint8_t v) { return MODE == 0 ? __popc((unsigned)(v & 7u)) : (int)(v & 0x7fu); }

template <int MODE>
__global__ void __launch_bounds__(kScanThreads) k_mc_scan_sums(const uint8_t* __restrict__ in, int64_t n, int32_t* __restrict__ sums) {
    __shared__ int32_t ws[kScanThreads / 32];
    const int64_t base = (int64_t)blockIdx.x * kScanTileMc + (int64_t)threadIdx.x * kScanPer;
    int32_t s = 0;
#pragma unroll
    for (int i = 0; i < kScanPer; ++i)
        if (base + i < n) s += mc_count_of<MODE>(in[base + i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        int32_t t = 0;
        for (int w = 0; w < kScanThreads / 32; ++w) t += ws[w];
        sums[blockIdx.x] = t;
    }
}

// one CTA: exclusive scan of the block sums in place, grand total -> *total
__global__ void __launch_bounds__(1024) k_mc_scan_top(int32_t* __restrict__ sums, int64_t nb, int32_t* __restrict__ total) {
    __shared__ int32_t ws[32];
    __shared__ int32_t carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int64_t b0 = 0; b0 < nb; b0 += 1024) {
        const int64_t i = b0 + threadIdx.x;
        const int32_t v = i < nb ? sums[i] : 0;
        int32_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if ((threadIdx.x & 31) >= o) incl += t;
        }
        if ((threadIdx.x & 31) == 31) ws[threadIdx.x >> 5] = incl;
        __syncthreads();
        if (threadIdx.x < 32) {
            const int32_t x = ws[threadIdx.x];
            int32_t wi = x;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int32_t t = __shfl_up_sync(0xffffffffu, wi, o);
                if ((int)threadIdx.x >= o) wi += t;
            }
            ws[threadIdx.x] = wi - x;
        }
        __syncthreads();
        const int32_t ex = carry + ws[threadIdx.x >> 5] + incl - v;
        if (i < nb) sums[i] = ex;
        __syncthreads();
        if (threadIdx.x == 1023) carry = ex + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) *total = carry;
}

template <int MODE>
__global__ void __launch_bounds__(kScanThreads) k_mc_scan_apply(const uint8_t* __restrict__ in, int64_t n, const int32_t* __restrict__ sums,
                                                                int32_t* __restrict__ out) {
    __shared__ int32_t ws[kScanThreads / 32];
    const int64_t base = (int64_t)blockIdx.x * kScanTileMc + (int64_t)threadIdx.x * kScanPer;
    int32_t v[kScanPer], s = 0;
#pragma unroll
    for (int i = 0; i < kScanPer; ++i) {
        v[i] = base + i < n ? mc_count_of<MODE>(in[base + i]) : 0;
        s += v[i];
    }
    int32_t incl = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if ((threadIdx.x & 31) >= o) incl += t;
    }
    if ((threadIdx.x & 31) == 31) ws[threadIdx.x >> 5] = incl;
    __syncthreads();
    int32_t wpre = 0;
    for (int w = 0; w < (int)(threadIdx.x >> 5); ++w) wpre += ws[w];
    int32_t run = sums[blockIdx.x] + wpre + incl - s;
#pragma unroll
    for (int i = 0; i < kScanPer; ++i) {
        if (base + i < n) out[base + i] = run;
        run += v[i];
    }
}

struct McEmit {
    const float* sdf;
    const uint8_t* ctri;
    const uint8_t* vflag;
    const int32_t* vbase;
    const int32_t* tbase;
    McGrid g;
    float level, ox, oy, oz, voxel;
    float* verts;
    int32_t* faces;
    int32_t vertex_base;
    int32_t x0;  // index of this slab's first x layer in the whole grid: coordinates are computed from GLOBAL grid indices
};

__global__ void k_mc_emit(const McEmit e) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= e.g.n) return;
    const uint8_t vf = e.vflag[p], ct = e.ctri[p];
    if (vf == 0 && (ct & 0x7fu) == 0) return;
    const int z = (int)(p % e.g.nz);
    const int64_t r = p / e.g.nz;
    const int y = (int)(r % e.g.ny), x = (int)(r / e.g.ny);
    if (vf) {  // vertices of the owned edges, in axis order
        const float v0 = e.sdf[p];
        int k = e.vbase[p];
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            if (!((vf >> a) & 1)) continue;
            int q3[3] = {x, y, z};
            q3[a] += 1;
            const float v1 = e.sdf[mc_lin(e.g, q3[0], q3[1], q3[2])];
            const float t = __fdiv_rn(__fsub_rn(e.level, v0), __fsub_rn(v1, v0));
            float gx = (float)(x + e.x0), gy = (float)y, gz = (float)z;
            if (a == 0) gx = __fadd_rn(gx, t);
            if (a == 1) gy = __fadd_rn(gy, t);
            if (a == 2) gz = __fadd_rn(gz, t);
            e.verts[3 * (int64_t)k] = __fadd_rn(__fmul_rn(gx, e.voxel), e.ox);
            e.verts[3 * (int64_t)k + 1] = __fadd_rn(__fmul_rn(gy, e.voxel), e.oy);
            e.verts[3 * (int64_t)k + 2] = __fadd_rn(__fmul_rn(gz, e.voxel), e.oz);
            ++k;
        }
    }
    const int nt = ct & 0x7f;
    if (nt) {
        int cs = 0;
#pragma unroll
        for (int c = 0; c < 8; ++c)
            cs |= (e.sdf[mc_lin(e.g, x + (c & 1), y + ((c >> 1) & 1), z + ((c >> 2) & 1))] < e.level ? 1 : 0) << c;
        int32_t* f = e.faces + 3 * (int64_t)e.tbase[p];
        for (int t = 0; t < 3 * nt; ++t) {
            const int ed = c_mc_tri[cs][t];
            const int a = ed >> 2, kk = ed & 3;
            const int o0 = a == 0 ? 1 : 0, o1 = a == 2 ? 1 : 2;
            int q3[3] = {x, y, z};
            q3[o0] += kk & 1;
            q3[o1] += kk >> 1;
            const int64_t q = mc_lin(e.g, q3[0], q3[1], q3[2]);
            f[t] = e.vertex_base + e.vbase[q] + __popc((unsigned)(e.vflag[q] & ((1u << a) - 1u)));
        }
    }
}

static inline int64_t mc_pad(int64_t b) { return ((b + 255) / 256) * 256; }
static inline int64_t mc_blocks(int64_t n) { return (n + kScanTileMc - 1) / kScanTileMc; }

}  // namespace pinq

using namespace pinq;

extern "C" {

// scratch: [ctri u8 n | vflag u8 n | vbase i32 n | tbase i32 n | block sums i32 2 * nb]
int64_t pinq_mc_scratch_bytes(int64_t n_points) {
    const int64_t n = n_points > 0 ? n_points : 1;
    return 2 * mc_pad(n) + 2 * mc_pad(4 * n) + 2 * mc_pad(4 * mc_blocks(n));
}

int pinq_mc_tables(int8_t* tri_out_256x16, uint8_t* ntri_out_256) {
    PINQ_REQUIRE(tri_out_256x16 && ntri_out_256, "null pointer");
    const McTables* T = mc_tables();
    if (!T) {
        set_error("marching-cubes case table could not be derived");
        return PINQ_ERR_CUDA;
    }
    memcpy(tri_out_256x16, T->tri, sizeof(T->tri));
    memcpy(ntri_out_256, T->ntri, sizeof(T->ntri));
    return PINQ_OK;
}

int pinq_mc_count(const float* sdf, const uint8_t* mask, const int32_t* dims3_host, float level, void* scratch, int32_t* d_counts2,
                  void* stream) {
    PINQ_REQUIRE(sdf && dims3_host && scratch && d_counts2, "null pointer");
    PINQ_REQUIRE(dims3_host[0] > 0 && dims3_host[1] > 0 && dims3_host[2] > 0, "empty grid");
    McGrid g{dims3_host[0], dims3_host[1], dims3_host[2], (int64_t)dims3_host[0] * dims3_host[1] * dims3_host[2]};
    PINQ_REQUIRE(g.n < (1ll << 31), "grid chunk too large for int32 offsets: split it into slabs");
    cudaStream_t st = as_stream(stream);
    int rc = mc_upload_tables(st);
    if (rc) return rc;
    unsigned char* s = reinterpret_cast<unsigned char*>(scratch);
    uint8_t* ctri = s;
    uint8_t* vflag = s + mc_pad(g.n);
    int32_t* vbase = reinterpret_cast<int32_t*>(s + 2 * mc_pad(g.n));
    int32_t* tbase = reinterpret_cast<int32_t*>(s + 2 * mc_pad(g.n) + mc_pad(4 * g.n));
    const int64_t nb = mc_blocks(g.n);
    int32_t* sums_v = reinterpret_cast<int32_t*>(s + 2 * mc_pad(g.n) + 2 * mc_pad(4 * g.n));
    int32_t* sums_t = sums_v + mc_pad(4 * nb) / 4;
    k_mc_cells<<<grid_for(g.n, 256), 256, 0, st>>>(sdf, mask, g, level, ctri);
    k_mc_edges<<<grid_for(g.n, 256), 256, 0, st>>>(sdf, ctri, g, level, vflag);
    k_mc_scan_sums<0><<<(unsigned)nb, kScanThreads, 0, st>>>(vflag, g.n, sums_v);
    k_mc_scan_top<<<1, 1024, 0, st>>>(sums_v, nb, d_counts2);
    k_mc_scan_apply<0><<<(unsigned)nb, kScanThreads, 0, st>>>(vflag, g.n, sums_v, vbase);
    k_mc_scan_sums<1><<<(unsigned)nb, kScanThreads, 0, st>>>(ctri, g.n, sums_t);
    k_mc_scan_top<<<1, 1024, 0, st>>>(sums_t, nb, d_counts2 + 1);
    k_mc_scan_apply<1><<<(unsigned)nb, kScanThreads, 0, st>>>(ctri, g.n, sums_t, tbase);
    return check_launch("pinq_mc_count");
}

int pinq_mc_emit(const float* sdf, const int32_t* dims3_host, int32_t x_first, float level, const float* origin3_host, float voxel,
                 const void* scratch, int32_t vertex_base, float* verts_out, int32_t* faces_out, void* stream) {
    PINQ_REQUIRE(sdf && dims3_host && origin3_host && scratch, "null pointer");
    McGrid g{dims3_host[0], dims3_host[1], dims3_host[2], (int64_t)dims3_host[0] * dims3_host[1] * dims3_host[2]};
    const unsigned char* s = reinterpret_cast<const unsigned char*>(scratch);
    McEmit e;
    e.sdf = sdf;
    e.ctri = s;
    e.vflag = s + mc_pad(g.n);
    e.vbase = reinterpret_cast<const int32_t*>(s + 2 * mc_pad(g.n));
    e.tbase = reinterpret_cast<const int32_t*>(s + 2 * mc_pad(g.n) + mc_pad(4 * g.n));
    e.g = g;
    e.level = level;
    e.ox = origin3_host[0]; e.oy = origin3_host[1]; e.oz = origin3_host[2];
    e.voxel = voxel;
    e.verts = verts_out;
    e.faces = faces_out;
    e.vertex_base = vertex_base;
    e.x0 = x_first;
    k_mc_emit<<<grid_for(g.n, 256), 256, 0, as_stream(stream)>>>(e);
    return check_launch("pinq_mc_emit");
}

}  // extern "C"

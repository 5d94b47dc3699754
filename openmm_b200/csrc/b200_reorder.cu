// b200_reorder.cu -- device-side spatial reordering of identical molecules along a Hilbert curve.
//
// Replaces ComputeContext::reorderAtomsImpl (platforms/common/src/ComputeContext.cpp:534-715), which downloads
// posq/velm, sorts on the host with libraries/hilbert and uploads again.  Here the whole operation stays on the
// GPU: molecule centres are wrapped into the box, binned on a 2^b Hilbert grid (counting sort: histogram,
// scan, scatter, per-bucket fix-up so the order never depends on atomic timing) and posq / posqCorrection /
// velm / atomIndex / posCellOffsets are permuted among the slots of identical molecules.  As in the reference,
// only molecules that are identical in every parameter are exchanged, so every index-based table (exclusion
// tiles, bonded terms, SETTLE clusters) stays valid.
#include "b200_internal.cuh"

__device__ __forceinline__ unsigned hilbertIndex3(unsigned x, unsigned y, unsigned z, int bits) {
    // Skilling's transpose algorithm (J. Skilling, "Programming the Hilbert curve", AIP Conf. Proc. 707, 2004)
    unsigned X[3] = {x, y, z};
    const unsigned M = 1u << (bits - 1);
    for (unsigned Q = M; Q > 1; Q >>= 1) {
        const unsigned P = Q - 1;
#pragma unroll
        for (int i = 0; i < 3; i++) {
            if (X[i] & Q) X[0] ^= P;
            else { const unsigned t = (X[0] ^ X[i]) & P; X[0] ^= t; X[i] ^= t; }
        }
    }
    X[1] ^= X[0]; X[2] ^= X[1];
    unsigned t = 0;
    for (unsigned Q = M; Q > 1; Q >>= 1) if (X[2] & Q) t ^= Q - 1;
    X[0] ^= t; X[1] ^= t; X[2] ^= t;
    unsigned h = 0;
    for (int b = bits - 1; b >= 0; b--) {
#pragma unroll
        for (int i = 0; i < 3; i++) h = (h << 1) | ((X[i] >> b) & 1u);
    }
    return h;
}

// wrap each molecule into the box (ComputeContext.cpp:605-636) and compute its bucket
__global__ void molWrapAndKey(int numMols, int atomsPerMol, const int* __restrict__ molBase, float4* posq, float4* posqCorr, int4* cellOffsets,
                              BoxD box, int bits, int* __restrict__ keys, int* __restrict__ count, int wrap) {
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= numMols) return;
    const int base = molBase[m];
    double cx = 0, cy = 0, cz = 0;
    for (int k = 0; k < atomsPerMol; k++) {
        const float4 p = posq[base + k], q = posqCorr[base + k];
        cx += (double)p.x + (double)q.x; cy += (double)p.y + (double)q.y; cz += (double)p.z + (double)q.z;
    }
    const double inv = 1.0 / atomsPerMol;
    cx *= inv; cy *= inv; cz *= inv;
    // (the centre is always reduced into the cell for the sort key; the atoms themselves move only in periodic systems)
    const int sz = (int)floor(cz * box.icz); cx -= sz * box.cx; cy -= sz * box.cy; cz -= sz * box.cz;
    const int sy = (int)floor(cy * box.iby); cx -= sy * box.bx; cy -= sy * box.by;
    const int sx = (int)floor(cx * box.iax); cx -= sx * box.ax;
    if (wrap && (sx != 0 || sy != 0 || sz != 0)) {
        const double tx = sx * box.ax + sy * box.bx + sz * box.cx, ty = sy * box.by + sz * box.cy, tz = sz * box.cz;
        for (int k = 0; k < atomsPerMol; k++) {
            const float4 p = posq[base + k], q = posqCorr[base + k];
            const double x = (double)p.x + (double)q.x - tx, y = (double)p.y + (double)q.y - ty, z = (double)p.z + (double)q.z - tz;
            posq[base + k] = make_float4((float)x, (float)y, (float)z, p.w);
            posqCorr[base + k] = make_float4((float)(x - (double)(float)x), (float)(y - (double)(float)y), (float)(z - (double)(float)z), 0.f);
            int4 o = cellOffsets[base + k]; o.x += sx; o.y += sy; o.z += sz; cellOffsets[base + k] = o;
        }
    }
    // fractional coordinates on the Hilbert grid
    const double fz = cz * box.icz, fy = (cy - fz * box.cy) * box.iby, fx = (cx - fy * box.bx - fz * box.cx) * box.iax;
    const int n = 1 << bits;
    const unsigned ix = (unsigned)min(n - 1, max(0, (int)(fx * n))), iy = (unsigned)min(n - 1, max(0, (int)(fy * n))), iz = (unsigned)min(n - 1, max(0, (int)(fz * n)));
    const int key = (int)hilbertIndex3(ix, iy, iz, bits);
    keys[m] = key;
    atomicAdd(&count[key], 1);
}

// ---- exclusive scan (cold path; three phases)
#define SCAN_BLOCK 1024
__global__ void scanBlocks(const int* __restrict__ in, int* __restrict__ out, int* __restrict__ blockSums, int n) {
    __shared__ int s[SCAN_BLOCK];
    const int i = blockIdx.x * SCAN_BLOCK + threadIdx.x;
    const int v = i < n ? in[i] : 0;
    s[threadIdx.x] = v;
    __syncthreads();
    for (int o = 1; o < SCAN_BLOCK; o <<= 1) {
        const int t = threadIdx.x >= o ? s[threadIdx.x - o] : 0;
        __syncthreads();
        s[threadIdx.x] += t;
        __syncthreads();
    }
    if (i < n) out[i] = s[threadIdx.x] - v;
    if (threadIdx.x == SCAN_BLOCK - 1) blockSums[blockIdx.x] = s[threadIdx.x];
}
__global__ void scanBlockSums(int* blockSums, int nb) {   // single block, serial over chunks
    __shared__ int s[SCAN_BLOCK];
    __shared__ int carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < nb; base += SCAN_BLOCK) {
        const int i = base + threadIdx.x;
        const int v = i < nb ? blockSums[i] : 0;
        s[threadIdx.x] = v;
        __syncthreads();
        for (int o = 1; o < SCAN_BLOCK; o <<= 1) {
            const int t = threadIdx.x >= o ? s[threadIdx.x - o] : 0;
            __syncthreads();
            s[threadIdx.x] += t;
            __syncthreads();
        }
        if (i < nb) blockSums[i] = s[threadIdx.x] - v + carry;
        __syncthreads();
        if (threadIdx.x == SCAN_BLOCK - 1) carry += s[threadIdx.x];
        __syncthreads();
    }
}
__global__ void scanAddOffsets(int* out, const int* __restrict__ blockSums, int n) {
    const int i = blockIdx.x * SCAN_BLOCK + threadIdx.x;
    if (i < n) out[i] += blockSums[blockIdx.x];
}

__global__ void molScatter(int numMols, const int* __restrict__ keys, const int* __restrict__ start, int* __restrict__ fill, int* __restrict__ order) {
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= numMols) return;
    const int k = keys[m];
    order[start[k] + atomicAdd(&fill[k], 1)] = m;
}
// make the order inside each bucket independent of atomic timing
__global__ void bucketFix(int numBuckets, int numMols, const int* __restrict__ start, const int* __restrict__ count, int* __restrict__ order) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= numBuckets) return;
    const int s = start[b], n = count[b];
    for (int i = 1; i < n; i++) {
        const int v = order[s + i];
        int j = i - 1;
        while (j >= 0 && order[s + j] > v) { order[s + j + 1] = order[s + j]; j--; }
        order[s + j + 1] = v;
    }
}
__global__ void molGather(int numMols, int atomsPerMol, const int* __restrict__ molBase, const int* __restrict__ order,
                          const float4* __restrict__ posqIn, const float4* __restrict__ corrIn, const double4* __restrict__ velIn, const int* __restrict__ idxIn, const int4* __restrict__ offIn,
                          float4* posq, float4* corr, double4* vel, int* idx, int4* off) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= numMols * atomsPerMol) return;
    const int p = t / atomsPerMol, k = t - p * atomsPerMol;
    const int dst = molBase[p] + k, src = molBase[order[p]] + k;
    posq[dst] = posqIn[src]; corr[dst] = corrIn[src]; vel[dst] = velIn[src]; idx[dst] = idxIn[src]; off[dst] = offIn[src];
}

void b200_reorder(b200_ctx* c) {
    ScopedTimer tm(c, B200_T_REORDER, c->stream);
    cudaStream_t s = c->stream;
    for (auto& g : c->groups) {
        const int nb = c->sortBuckets;
        const int grid = (g.numMols + 255) / 256;
        CUDA_CHECK(cudaMemsetAsync(c->sortCount, 0, sizeof(int) * (size_t)nb, s));
        molWrapAndKey<<<grid, 256, 0, s>>>(g.numMols, g.atomsPerMol, g.dBase, c->posq, c->posqCorr, c->cellOffsets, c->boxD, c->sortBits, c->sortKeys, c->sortCount,
                                          (c->hasNonbonded && (c->nb.method == B200_NB_NOCUTOFF || c->nb.method == B200_NB_CUTOFF_NONPERIODIC)) ? 0 : 1);
        LAUNCH_COUNT(c, 1);
        if (g.numMols < 2) continue;   // unique molecules are only wrapped into the box
        const int nblk = (nb + SCAN_BLOCK - 1) / SCAN_BLOCK;
        int* blockSums = c->sortStart + nb;   // tail of the allocation
        scanBlocks<<<nblk, SCAN_BLOCK, 0, s>>>(c->sortCount, c->sortStart, blockSums, nb);
        scanBlockSums<<<1, SCAN_BLOCK, 0, s>>>(blockSums, nblk);
        scanAddOffsets<<<nblk, SCAN_BLOCK, 0, s>>>(c->sortStart, blockSums, nb);
        int* fill = c->sortStart + nb + nblk + 8;
        CUDA_CHECK(cudaMemsetAsync(fill, 0, sizeof(int) * (size_t)nb, s));
        molScatter<<<grid, 256, 0, s>>>(g.numMols, c->sortKeys, c->sortStart, fill, c->sortOrder);
        bucketFix<<<(nb + 255) / 256, 256, 0, s>>>(nb, g.numMols, c->sortStart, c->sortCount, c->sortOrder);
        // permute: copy the live arrays aside, then gather back
        CUDA_CHECK(cudaMemcpyAsync(c->posqTmp, c->posq, sizeof(float4) * c->Npad, cudaMemcpyDeviceToDevice, s));
        CUDA_CHECK(cudaMemcpyAsync(c->posqCorrTmp, c->posqCorr, sizeof(float4) * c->Npad, cudaMemcpyDeviceToDevice, s));
        CUDA_CHECK(cudaMemcpyAsync(c->velmTmp, c->velm, sizeof(double4) * c->Npad, cudaMemcpyDeviceToDevice, s));
        CUDA_CHECK(cudaMemcpyAsync(c->atomIndexTmp, c->atomIndex, sizeof(int) * c->Npad, cudaMemcpyDeviceToDevice, s));
        CUDA_CHECK(cudaMemcpyAsync(c->cellOffsetsTmp, c->cellOffsets, sizeof(int4) * c->Npad, cudaMemcpyDeviceToDevice, s));
        const int total = g.numMols * g.atomsPerMol;
        molGather<<<(total + 255) / 256, 256, 0, s>>>(g.numMols, g.atomsPerMol, g.dBase, c->sortOrder,
            c->posqTmp, c->posqCorrTmp, c->velmTmp, c->atomIndexTmp, c->cellOffsetsTmp, c->posq, c->posqCorr, c->velm, c->atomIndex, c->cellOffsets);
        LAUNCH_COUNT(c, 7);
    }
    // molecules that are unique are still wrapped into the box
    CUDA_CHECK(cudaGetLastError());
    c->hostOrderStale = true;
    c->stepsSinceReorder = 0;
    b200_nblist_force_rebuild(c);
}

// Bag-of-words transform: Frame::ComputeBoW -> DBoW3::Vocabulary::transform(features, BowVector, FeatureVector, levelsup)
// (src/datastructures/frame.cpp:455-462; thirdparty/DBoW3/DBoW3/src/Vocabulary.cpp:754-876, BowVector.cpp:35-88,
// FeatureVector.cpp:31-45, DescManip.cpp:92-119).
//
//   k_bow_walk   one warp per descriptor walks the vocabulary tree: at every level each lane takes one child of the
//                current node (32 bytes = one sector of the node-descriptor table, which stays L2-resident: ORBvoc is
//                1.1 M nodes x 32 B = 35 MB), computes the Hamming distance and the warp takes the minimum of
//                (distance << 22 | child position) - DBoW3's "first strictly smaller distance wins" rule.
//   k_bow_build  one CTA per frame turns the per-descriptor (word, node) pairs into the two std::map results: a
//                bitonic sort of (id << 32 | feature index) in shared memory, run heads -> entries, weights summed
//                in the reference's order (feature order inside a word, ascending word id for the norm) so the
//                doubles come out bit-identical.
#pragma once
#include "kernels_match.cuh"

struct VocabDev {
    int nnodes, levels, weighting, scoring, anyWord;
    const int* childStart;      // [nnodes + 1]
    const int* children;        // node ids, DBoW3's Node::children order
    const uint4* desc;          // [nnodes][2]
    const double* weight;       // [nnodes]
    const int* wordId;          // [nnodes], -1 for inner nodes
};

#define VIORB_BOW_W_TF_IDF 0
#define VIORB_BOW_W_TF 1
#define VIORB_BOW_S_L2 1
#define VIORB_BOW_S_DOT 5

// grid (ceil(cap / 8), nframes), block 256.  leaf / nid: [nframes][cap].
__global__ void __launch_bounds__(256) k_bow_walk(VocabDev v, const uint8_t* __restrict__ desc, const int* __restrict__ counts, int cap,
                                                  int levelsup, int* __restrict__ leaf, int* __restrict__ nid) {
    const int lane = threadIdx.x & 31;
    const int f = blockIdx.y;
    const int i = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (i >= min(counts[f], cap)) return;
    const uint4* q = reinterpret_cast<const uint4*>(desc + ((size_t)f * cap + i) * 32);
    const uint4 q0 = __ldg(q), q1 = __ldg(q + 1);
    const int nidLevel = v.levels - levelsup;
    int node = 0, level = 0, nodeAtLevel = 0;
    int cs = __ldg(v.childStart), ce = __ldg(v.childStart + 1);
    while (ce > cs) {                                   // do { ... } while (!isLeaf()): the root is entered unconditionally,
        ++level;                                        // and a childless root leaves the loop with final_id = 0 either way
        unsigned best = 0xFFFFFFFFu;
        for (int c0 = cs; c0 < ce; c0 += 32) {
            const int c = c0 + lane;
            unsigned key = 0xFFFFFFFFu;
            if (c < ce) {
                const int id = __ldg(v.children + c);
                const uint4 b0 = __ldg(v.desc + (size_t)id * 2), b1 = __ldg(v.desc + (size_t)id * 2 + 1);
                const int2 h = hamming256_csa(q0, q1, b0, b1);
                key = ((unsigned)(h.x + 2 * h.y) << 22) | (unsigned)(c - cs);
            }
            best = min(best, __reduce_min_sync(0xffffffffu, key));
        }
        node = __ldg(v.children + cs + (int)(best & 0x3FFFFFu));
        if (level == nidLevel) nodeAtLevel = node;
        cs = __ldg(v.childStart + node);
        ce = __ldg(v.childStart + node + 1);
    }
    if (lane == 0) {
        leaf[(size_t)f * cap + i] = node;
        nid[(size_t)f * cap + i] = nodeAtLevel;
    }
}

// In-place ascending bitonic sort of P (power of two) 64-bit keys in shared memory.
__device__ __forceinline__ void bow_bitonic_sort(unsigned long long* keys, int P) {
    for (int k = 2; k <= P; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = threadIdx.x; t < P; t += blockDim.x) {
                const int p = t ^ j;
                if (p > t) {
                    const unsigned long long a = keys[t], b = keys[p];
                    const bool up = (t & k) == 0;
                    if ((a > b) == up) { keys[t] = b; keys[p] = a; }
                }
            }
            __syncthreads();
        }
    }
}

// Exclusive block scan of one flag per sorted slot, done in chunks of blockDim.x; heads[] receives, for every
// run head, its rank.  Returns the number of heads.  scratch: blockDim.x / 32 + 1 ints.
__device__ __forceinline__ int bow_block_count_before(int flag, int* scratch, int& total) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const unsigned m = __ballot_sync(0xffffffffu, flag != 0);
    if (lane == 0) scratch[wid] = __popc(m);
    __syncthreads();
    int before = __popc(m & ((1u << lane) - 1));
    int sum = 0;
    for (int w = 0; w < nw; ++w) {
        const int c = scratch[w];
        if (w < wid) before += c;
        sum += c;
    }
    total = sum;
    __syncthreads();
    return before;
}

// grid (nframes), block 256, dynamic shared memory P * 8 bytes (P = power of two >= cap).
// Outputs per frame f (row stride cap; fvStart cap + 1): bowWord / bowValue [nBow[f]], fvNode / fvStart [nFv[f] (+1)],
// fvFeature [fvStart[nFv[f]]].
__global__ void __launch_bounds__(256) k_bow_build(VocabDev v, const int* __restrict__ counts, int cap, int P, const int* __restrict__ leaf,
                                                   const int* __restrict__ nid, uint32_t* __restrict__ bowWord,
                                                   double* __restrict__ bowValue, int* __restrict__ nBow, uint32_t* __restrict__ fvNode,
                                                   int* __restrict__ fvStart, uint32_t* __restrict__ fvFeature, int* __restrict__ nFv) {
    VIORB_DYN_SMEM(smemRaw);
    unsigned long long* keys = reinterpret_cast<unsigned long long*>(smemRaw);
    __shared__ int scratch[9];
    __shared__ int shBase;
    __shared__ double shNorm;
    const int f = blockIdx.x, tid = threadIdx.x;
    const int n = min(counts[f], cap);
    const int* lf = leaf + (size_t)f * cap;
    const int* nd = nid + (size_t)f * cap;
    uint32_t* bw = bowWord + (size_t)f * cap;
    double* bv = bowValue + (size_t)f * cap;
    uint32_t* fn = fvNode + (size_t)f * cap;
    int* fs = fvStart + (size_t)f * (cap + 1);
    uint32_t* ff = fvFeature + (size_t)f * cap;
    if (!v.anyWord || n == 0) {                         // empty vocabulary (Vocabulary.cpp:761) or no features
        if (tid == 0) { nBow[f] = 0; nFv[f] = 0; fs[0] = 0; }
        return;
    }
    const bool tf = v.weighting == VIORB_BOW_W_TF_IDF || v.weighting == VIORB_BOW_W_TF;
    const bool must = v.scoring != VIORB_BOW_S_DOT;     // ScoringObject.h:72-91
    const bool l2 = v.scoring == VIORB_BOW_S_L2;
    const unsigned long long EMPTY = ~0ull;

    // ---- BowVector: sort (word id, feature index); stopped words (weight <= 0) drop out ----
    for (int t = tid; t < P; t += blockDim.x) {
        unsigned long long k = EMPTY;
        if (t < n) {
            const int l = lf[t];
            if (v.weight[l] > 0) k = ((unsigned long long)(unsigned)v.wordId[l] << 32) | (unsigned)t;
        }
        keys[t] = k;
    }
    if (tid == 0) shBase = 0;
    __syncthreads();
    bow_bitonic_sort(keys, P);
    for (int base = 0; base < P; base += blockDim.x) {
        const int t = base + tid;
        const unsigned long long k = keys[t];
        const bool live = k != EMPTY;
        const bool head = live && (t == 0 || (unsigned)(keys[t - 1] >> 32) != (unsigned)(k >> 32));
        int total;
        const int before = bow_block_count_before(head, scratch, total);
        if (head) {
            const int pos = shBase + before;
            const unsigned word = (unsigned)(k >> 32);
            const double w = v.weight[lf[(unsigned)k]];
            double acc = w;                               // BowVector::addWeight: insert w, then += w per further feature
            if (tf)
                for (int u = t + 1; u < P && (unsigned)(keys[u] >> 32) == word && keys[u] != EMPTY; ++u) acc = __dadd_rn(acc, w);
            bw[pos] = word;
            bv[pos] = acc;
        }
        __syncthreads();
        if (tid == 0) shBase += total;
        __syncthreads();
    }
    const int nb = shBase;
    if (tid == 0) {
        // serial on purpose: the reference sums in std::map order and double addition is not associative
        double norm = 0.0;
        if (must) {
            if (l2) {
                for (int k = 0; k < nb; ++k) norm = __dadd_rn(norm, __dmul_rn(bv[k], bv[k]));
                norm = __dsqrt_rn(norm);
            } else {
                for (int k = 0; k < nb; ++k) norm = __dadd_rn(norm, fabs(bv[k]));
            }
        } else if (tf) {
            norm = (double)nb;                            // "unnecessary when normalizing": v /= v.size()
        }
        shNorm = norm;
        nBow[f] = nb;
    }
    __syncthreads();
    {
        const double norm = shNorm;
        if (norm > 0.0)
            for (int k = tid; k < nb; k += blockDim.x) bv[k] = __ddiv_rn(bv[k], norm);
    }
    __syncthreads();

    // ---- FeatureVector: sort (node id at level L - levelsup, feature index) over the same live features ----
    for (int t = tid; t < P; t += blockDim.x) {
        unsigned long long k = EMPTY;
        if (t < n && v.weight[lf[t]] > 0) k = ((unsigned long long)(unsigned)nd[t] << 32) | (unsigned)t;
        keys[t] = k;
    }
    if (tid == 0) shBase = 0;
    __syncthreads();
    bow_bitonic_sort(keys, P);
    int liveTotal = 0;
    for (int base = 0; base < P; base += blockDim.x) {
        const int t = base + tid;
        const unsigned long long k = keys[t];
        const bool live = k != EMPTY;
        const bool head = live && (t == 0 || (unsigned)(keys[t - 1] >> 32) != (unsigned)(k >> 32));
        if (live) ff[t] = (unsigned)k;                    // live keys sort before EMPTY: slot t is the CSR position
        int total;
        const int before = bow_block_count_before(head, scratch, total);
        if (head) {
            fn[shBase + before] = (unsigned)(k >> 32);
            fs[shBase + before] = t;
        }
        int liveHere;
        bow_block_count_before(live, scratch, liveHere);
        liveTotal += liveHere;
        if (tid == 0) shBase += total;
        __syncthreads();
    }
    if (tid == 0) { nFv[f] = shBase; fs[shBase] = liveTotal; }
}

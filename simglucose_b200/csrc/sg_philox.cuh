// Counter-based Philox4x32-10 stream shared by the generator kernels (sg_generators.cu) and the random-action
// controller of the episode kernels (sg_kernels.cu): key = seed, counter = (draw block, substream tag, 64-bit stream id).
// Any shard of a population can therefore draw its streams independently on any GPU.
#pragma once
#include <cstdint>

namespace sgp {

struct Philox {
    uint32_t key[2];
    uint32_t ctr[4];
    uint32_t out[4];
    int have;

    // block: first value of the block counter (a caller that needs random access into a stream -- the action of
    // env-step k -- starts at block k; sequential callers start at 0)
    __device__ Philox(uint64_t seed, uint64_t stream, uint32_t substream, uint32_t block = 0)
    {
        key[0] = (uint32_t)seed; key[1] = (uint32_t)(seed >> 32);
        ctr[0] = block; ctr[1] = substream; ctr[2] = (uint32_t)stream; ctr[3] = (uint32_t)(stream >> 32);
        have = 0;
    }
    __device__ void round(uint32_t (&c)[4], const uint32_t (&k)[2])
    {
        const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
        const uint32_t hi0 = __umulhi(M0, c[0]), lo0 = M0 * c[0];
        const uint32_t hi1 = __umulhi(M1, c[2]), lo1 = M1 * c[2];
        const uint32_t n0 = hi1 ^ c[1] ^ k[0], n1 = lo1, n2 = hi0 ^ c[3] ^ k[1], n3 = lo0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
    }
    __device__ void refill()
    {
        uint32_t c[4] = {ctr[0], ctr[1], ctr[2], ctr[3]};
        uint32_t k[2] = {key[0], key[1]};
#pragma unroll
        for (int r = 0; r < 10; r++) {
            round(c, k);
            k[0] += 0x9E3779B9u; k[1] += 0xBB67AE85u;
        }
        out[0] = c[0]; out[1] = c[1]; out[2] = c[2]; out[3] = c[3];
        ctr[0]++;
        have = 4;
    }
    // Random access: the four words of draw block `blk` of this stream (what refill() leaves in out[] when ctr[0] == blk).
    __device__ void block_at(uint32_t blk, uint32_t (&w)[4]) const
    {
        uint32_t c[4] = {blk, ctr[1], ctr[2], ctr[3]};
        uint32_t k[2] = {key[0], key[1]};
#pragma unroll
        for (int r = 0; r < 10; r++) {
            const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
            const uint32_t hi0 = __umulhi(M0, c[0]), lo0 = M0 * c[0];
            const uint32_t hi1 = __umulhi(M1, c[2]), lo1 = M1 * c[2];
            const uint32_t n0 = hi1 ^ c[1] ^ k[0], n1 = lo1, n2 = hi0 ^ c[3] ^ k[1], n3 = lo0;
            c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
            k[0] += 0x9E3779B9u; k[1] += 0xBB67AE85u;
        }
        w[0] = c[0]; w[1] = c[1]; w[2] = c[2]; w[3] = c[3];
    }
    // The value uniform() builds from the words (a, b) = (next32(), next32()).
    static __device__ double uniform_of(uint32_t a, uint32_t b)
    {
        const uint64_t bits = (((uint64_t)a << 32) | b) >> 11;
        return ((double)bits + 0.5) * (1.0 / 9007199254740992.0);
    }
    // The j-th value a sequential caller of normal() (and nothing else) sees: draw block j / 2, words (3, 2) for even j
    // and (1, 0) for odd j -- next32() hands the words out from the top.
    static __device__ double normal_of(const uint32_t (&w)[4], int odd)
    {
        return normcdfinv(odd ? uniform_of(w[1], w[0]) : uniform_of(w[3], w[2]));
    }
    __device__ uint32_t next32()
    {
        if (have == 0) refill();
        return out[--have];
    }
    // uniform in (0, 1), 53 bits
    __device__ double uniform()
    {
        const uint32_t a = next32(), b = next32();
        return uniform_of(a, b);
    }
    __device__ double normal()
    {
        // inverse-CDF normal: one uniform per draw, monotone, no rejection
        return normcdfinv(uniform());
    }
};

}  // namespace sgp

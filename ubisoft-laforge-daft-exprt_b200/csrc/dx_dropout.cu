// Export of the dropout masks of the training step (dx_dropout.cuh) for the oracle.
#include "dx_common.cuh"
#include "dx_ptx.cuh"
#include "dx_dropout.cuh"

namespace dx {

// mask export (uint8 keep flags) in the reference's padded coordinates: rows kind -> [B][n_max][C]; attention -> [B][2][n_max][n_max]
__global__ void dropout_mask_kernel(uint8_t* __restrict__ out, int attention, int B, int n_max, int C, DropSpec d, uint32_t site, int64_t n) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t z;
    int e;
    if (!attention) {
        const int c = (int)(i % C);
        const int64_t bp = i / C;
        z = drop_hash(d.seed, site, drop_pair_rows((int)(bp / n_max), (int)(bp % n_max), n_max, C, c & ~1));
        e = c & 1;
    } else {
        const int k = (int)(i % n_max);
        const int64_t bhq = i / n_max;
        const int q = (int)(bhq % n_max), h = (int)((bhq / n_max) % 2), b = (int)(bhq / n_max / 2);
        z = drop_hash(d.seed, site, drop_pair_attn_row(b, h, q, n_max) + (k >> 1));
        e = k & 1;
    }
    out[i] = drop_keep(z, e, d.thr) ? 1 : 0;
}

}  // namespace dx

using namespace dx;

extern "C" int dx_dropout_mask(uint64_t seed, int kind, int stack, int block, float p, int B, int n_max, int C, uint8_t* out,
                               void* stream) {
    DX_REQUIRE(kind >= 0 && kind <= 3 && stack >= 0 && stack <= 2 && block >= 0 && block <= 3, "dx_dropout_mask: bad site");
    const int attention = kind == (int)DROP_ATTN_W;
    DX_REQUIRE(out != nullptr && B >= 1 && n_max >= 1 && (attention || (C >= 2 && C % 2 == 0)), "dx_dropout_mask: bad arguments");
    const int64_t n = attention ? (int64_t)B * 2 * n_max * n_max : (int64_t)B * n_max * C;
    const DropSpec d = make_drop(seed, p);
    dropout_mask_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(out, attention, B, n_max, C, d, drop_site(kind, stack, block), n);
    DX_CHECK_LAUNCH();
    return 0;
}

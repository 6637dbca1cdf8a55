// Host-side batch collation for bulk generation (BASELINE config 3): the padded-batch assembly of
// the reference's DaftExprtDataCollate.__call__ (data_loader.py:214-263) / the fine_tune loop (fine_tune.py:67-80) as
// plain memcpy / memset into ONE pinned staging block, in native code so that it runs without the Python GIL and on
// several host threads.  No CUDA calls: the block crosses PCIe as one cudaMemcpyAsync issued by the caller.
#include <string.h>
#include <thread>
#include <vector>
#include "dx_common.cuh"

namespace {

// forward-tuple order of the offsets (model.py:893-896): symbols, durations_float, durations_int, symbols_energy,
// symbols_pitch, input_lengths, frames_energy, frames_pitch, mel_specs, output_lengths, speaker_ids, spk_embs
enum { O_SYM, O_DFL, O_DINT, O_SEN, O_SPI, O_ILEN, O_FEN, O_FPI, O_MEL, O_OLEN, O_SPK, O_EMB };
// per-utterance source fields
enum { F_SYM, F_DFL, F_DINT, F_SEN, F_SPI, F_FEN, F_FPI, F_MEL, F_SPK, F_EMB, F_COUNT };

inline void copy_pad(uint8_t* dst, const void* src, size_t n_bytes, size_t row_bytes) {
    memcpy(dst, src, n_bytes);
    if (row_bytes > n_bytes) memset(dst + n_bytes, 0, row_bytes - n_bytes);
}

// fs == nullptr: padded frame side ((B, T) / (B, M, T)).  fs != nullptr (DX_IO_PACKED_FRAMES): utterance b's frames start at
// fs[b]; energy / pitch are float[sum T_b] and the mel is ONE contiguous (M, T_b) block per utterance at M * fs[b] - the
// source array as it is, so the whole utterance is a single memcpy and nothing is padded.
void collate_range(int b0, int b1, int L, int T, int M, int E, const void* const* ptrs, const int32_t* lens, uint8_t* block,
                   const int64_t* offs, const int64_t* fs) {
    for (int b = b0; b < b1; ++b) {
        const void* const* p = ptrs + (size_t)b * F_COUNT;
        const size_t l = (size_t)lens[2 * b], t = (size_t)lens[2 * b + 1];
        if (fs != nullptr) {
            memcpy(block + offs[O_FEN] + (size_t)fs[b] * 4, p[F_FEN], t * 4);
            memcpy(block + offs[O_FPI] + (size_t)fs[b] * 4, p[F_FPI], t * 4);
            memcpy(block + offs[O_MEL] + (size_t)fs[b] * M * 4, p[F_MEL], t * M * 4);
        } else {
            copy_pad(block + offs[O_FEN] + (size_t)b * T * 4, p[F_FEN], t * 4, (size_t)T * 4);
            copy_pad(block + offs[O_FPI] + (size_t)b * T * 4, p[F_FPI], t * 4, (size_t)T * 4);
            const float* mel = static_cast<const float*>(p[F_MEL]);
            uint8_t* dmel = block + offs[O_MEL] + (size_t)b * M * T * 4;
            for (int m = 0; m < M; ++m) copy_pad(dmel + (size_t)m * T * 4, mel + (size_t)m * t, t * 4, (size_t)T * 4);
        }
        copy_pad(block + offs[O_SYM] + (size_t)b * L * 8, p[F_SYM], l * 8, (size_t)L * 8);
        copy_pad(block + offs[O_DINT] + (size_t)b * L * 8, p[F_DINT], l * 8, (size_t)L * 8);
        copy_pad(block + offs[O_DFL] + (size_t)b * L * 4, p[F_DFL], l * 4, (size_t)L * 4);
        copy_pad(block + offs[O_SEN] + (size_t)b * L * 4, p[F_SEN], l * 4, (size_t)L * 4);
        copy_pad(block + offs[O_SPI] + (size_t)b * L * 4, p[F_SPI], l * 4, (size_t)L * 4);
        reinterpret_cast<int64_t*>(block + offs[O_ILEN])[b] = (int64_t)l;
        reinterpret_cast<int64_t*>(block + offs[O_OLEN])[b] = (int64_t)t;
        reinterpret_cast<int64_t*>(block + offs[O_SPK])[b] = *static_cast<const int64_t*>(p[F_SPK]);
        memcpy(block + offs[O_EMB] + (size_t)b * E * 4, p[F_EMB], (size_t)E * 4);
    }
}

}  // namespace

static int host_collate(int B, int L, int T, int M, int E, const void* const* ptrs, const int32_t* lens, void* block,
                        const int64_t* offs, int threads, bool packed) {
    DX_REQUIRE(B >= 1 && L >= 1 && T >= 1 && M >= 1 && E >= 1 && ptrs && lens && block && offs, "dx_host_collate: bad arguments");
    for (int b = 0; b < B; ++b) {
        DX_REQUIRE(lens[2 * b] >= 0 && lens[2 * b] <= L && lens[2 * b + 1] >= 0 && lens[2 * b + 1] <= T,
                   "dx_host_collate: utterance %d has %d symbols / %d frames, batch is padded to %d / %d", b, lens[2 * b], lens[2 * b + 1], L, T);
        for (int f = 0; f < F_COUNT; ++f) DX_REQUIRE(ptrs[(size_t)b * F_COUNT + f] != nullptr, "dx_host_collate: NULL field %d of utterance %d", f, b);
    }
    uint8_t* blk = static_cast<uint8_t*>(block);
    std::vector<int64_t> fs_store;
    const int64_t* fs = nullptr;
    if (packed) {
        fs_store.resize((size_t)B + 1);
        fs_store[0] = 0;
        for (int b = 0; b < B; ++b) fs_store[b + 1] = fs_store[b] + lens[2 * b + 1];
        fs = fs_store.data();
    }
    if (threads > B) threads = B;
    if (threads <= 1) {
        collate_range(0, B, L, T, M, E, ptrs, lens, blk, offs, fs);
        return 0;
    }
    std::vector<std::thread> pool;
    pool.reserve(threads - 1);
    const int per = (B + threads - 1) / threads;
    for (int i = 1; i < threads; ++i) {
        const int b0 = i * per, b1 = b0 + per < B ? b0 + per : B;
        if (b0 < b1) pool.emplace_back(collate_range, b0, b1, L, T, M, E, ptrs, lens, blk, offs, fs);
    }
    collate_range(0, per < B ? per : B, L, T, M, E, ptrs, lens, blk, offs, fs);
    for (auto& th : pool) th.join();
    return 0;
}

extern "C" int dx_host_collate(int B, int L, int T, int M, int E, const void* const* ptrs, const int32_t* lens, void* block,
                               const int64_t* offs, int threads) {
    return host_collate(B, L, T, M, E, ptrs, lens, block, offs, threads, false);
}

extern "C" int dx_host_collate_packed(int B, int L, int T, int M, int E, const void* const* ptrs, const int32_t* lens,
                                      void* block, const int64_t* offs, int threads) {
    return host_collate(B, L, T, M, E, ptrs, lens, block, offs, threads, true);
}

// Host-side staging helper (no device work): packs one batch of the TF-form feed -- numeric weights, per-field LOCAL
// categorical ids, categorical weights (FNN.py:41-43; the graph adds the field offsets, FNN.py:65-69) -- into the
// [N, F] int32 ids / float32 wts blocks of the pinned staging buffer, rows split over a few threads.  numpy needs
// ~6 ms for this at N = 65536, F = 39 (strided int64 -> int32 writes); the device step takes 2 ms.
#include <thread>
#include <vector>

#include "common.cuh"

namespace {

template <class IdT>
void pack_rows(const IdT* c_ids, const int64_t* offsets, const float* v_wts, const float* c_wts, int64_t r0, int64_t r1,
               int n_num, int n_cat, int32_t* ids, float* wts) {
    const int F = n_num + n_cat;
    for (int64_t n = r0; n < r1; ++n) {
        int32_t* id = ids + n * F;
        float* w = wts + n * F;
        for (int j = 0; j < n_num; ++j) { id[j] = j; w[j] = v_wts[n * n_num + j]; }
        const IdT* c = c_ids + n * n_cat;
        for (int j = 0; j < n_cat; ++j) id[n_num + j] = (int32_t)((int64_t)c[j] + offsets[j]);
        if (c_wts) {
            const float* cw = c_wts + n * n_cat;
            for (int j = 0; j < n_cat; ++j) w[n_num + j] = cw[j];
        } else {
            for (int j = 0; j < n_cat; ++j) w[n_num + j] = 1.f;
        }
    }
}

}  // namespace

extern "C" int dctr_host_pack_fields(const void* c_ids, int id_bytes, const int64_t* offsets, const float* v_wts,
                                     const float* c_wts, int64_t N, int n_num, int n_cat, int32_t* ids, float* wts,
                                     int threads) {
    using namespace dctr;
    DCTR_REQUIRE(N >= 0 && n_num >= 0 && n_cat >= 0 && n_num + n_cat > 0, "host_pack_fields: bad shape");
    DCTR_REQUIRE(id_bytes == 4 || id_bytes == 8, "host_pack_fields: ids must be int32 or int64");
    DCTR_REQUIRE(ids && wts && (n_cat == 0 || (c_ids && offsets)) && (n_num == 0 || v_wts), "host_pack_fields: null pointer");
    if (N == 0) return DCTR_OK;
    if (threads < 1) threads = 1;
    if (threads > 16) threads = 16;
    const int64_t min_rows = 4096;                      // below this a thread start costs more than the copy
    int T = (int)((N + min_rows - 1) / min_rows);
    if (T > threads) T = threads;
    auto work = [&](int t) {
        const int64_t r0 = N * t / T, r1 = N * (t + 1) / T;
        if (id_bytes == 8)
            pack_rows(static_cast<const int64_t*>(c_ids), offsets, v_wts, c_wts, r0, r1, n_num, n_cat, ids, wts);
        else
            pack_rows(static_cast<const int32_t*>(c_ids), offsets, v_wts, c_wts, r0, r1, n_num, n_cat, ids, wts);
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < T; ++t) pool.emplace_back(work, t);
    work(0);
    for (auto& th : pool) th.join();
    return DCTR_OK;
}

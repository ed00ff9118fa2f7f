// shard.cpp -- multi-GPU decode inside one process: frames are independent, so a batch is cut into
// contiguous frame ranges, one per device, each decoded by its own instance on its own host thread. There is no
// exchange step and therefore no collective; the host "concatenates" by letting every device write its slice of
// the caller's output buffer. (bench.py's multi-GPU runs use one process per GPU instead and shard the same way.)
#include "../../include/b200_viterbi.h"

#include <cstdio>
#include <cstring>
#include <new>
#include <string>
#include <thread>
#include <vector>

extern "C" int b200v_shard_range(int n_frames, int n_shards, int shard, int* frame0, int* count) {
    if (n_frames < 0 || n_shards <= 0 || shard < 0 || shard >= n_shards || !frame0 || !count) return B200V_ERR_ARG;
    // contiguous ranges [g*B/G, (g+1)*B/G): sizes differ by at most one frame
    const long long lo = (long long)shard * n_frames / n_shards, hi = (long long)(shard + 1) * n_frames / n_shards;
    *frame0 = (int)lo;
    *count = (int)(hi - lo);
    return B200V_OK;
}

// Shards proportional to weights (e.g. the host-to-device rate each GPU's link sustains when all of them copy at once,
// tools/h2d_probe.py): boundary g = round(B * prefix_weight(g) / total_weight), so the ranges are contiguous, cover every
// frame once and depend only on the weights. Equal weights give b200v_shard_range's ranges up to rounding.
extern "C" int b200v_shard_range_weighted(int n_frames, const double* weights, int n_shards, int shard, int* frame0, int* count) {
    if (n_frames < 0 || n_shards <= 0 || shard < 0 || shard >= n_shards || !frame0 || !count || !weights) return B200V_ERR_ARG;
    double total = 0, before = 0;
    for (int g = 0; g < n_shards; g++) {
        if (!(weights[g] >= 0)) return B200V_ERR_ARG;       // negative or NaN
        total += weights[g];
        if (g < shard) before += weights[g];
    }
    if (!(total > 0)) return B200V_ERR_ARG;
    const long long lo = shard == 0 ? 0 : (long long)((double)n_frames * before / total + 0.5);
    const long long hi = shard == n_shards - 1 ? n_frames : (long long)((double)n_frames * (before + weights[shard]) / total + 0.5);
    *frame0 = (int)lo;
    *count = (int)(hi > lo ? hi - lo : 0);
    return B200V_OK;
}

struct b200v_sharded {
    int n_frames = 0, frame_steps = 0, sym_row = 0;
    std::vector<b200v_batch*> inst;
    std::vector<int> frame0, count, device;
    std::string error;
};

extern "C" b200v_sharded* b200v_sharded_create(int codec, int K, int R, const int* poly, int frame_steps, int n_frames,
                                               const int* devices, int n_devices, const b200v_wy_config* wy_config) {
    return b200v_sharded_create_weighted(codec, K, R, poly, frame_steps, n_frames, devices, nullptr, n_devices, wy_config);
}

extern "C" b200v_sharded* b200v_sharded_create_weighted(int codec, int K, int R, const int* poly, int frame_steps, int n_frames,
                                                        const int* devices, const double* weights, int n_devices,
                                                        const b200v_wy_config* wy_config) {
    if (!devices || n_devices <= 0 || n_frames <= 0) return nullptr;
    b200v_sharded* s = new (std::nothrow) b200v_sharded();
    if (!s) return nullptr;
    s->n_frames = n_frames; s->frame_steps = frame_steps;
    for (int g = 0; g < n_devices; g++) {
        int f0 = 0, cnt = 0;
        const int rc = weights ? b200v_shard_range_weighted(n_frames, weights, n_devices, g, &f0, &cnt)
                               : b200v_shard_range(n_frames, n_devices, g, &f0, &cnt);
        if (rc != B200V_OK) {
            for (b200v_batch* p : s->inst) b200v_batch_destroy(p);
            delete s;
            return nullptr;
        }
        b200v_batch* h = cnt > 0 ? b200v_batch_create(codec, K, R, poly, frame_steps, cnt, devices[g], wy_config) : nullptr;
        if (cnt > 0 && !h) {
            for (b200v_batch* p : s->inst) b200v_batch_destroy(p);
            delete s;
            return nullptr;           // b200v_last_error() holds the reason
        }
        s->inst.push_back(h); s->frame0.push_back(f0); s->count.push_back(cnt); s->device.push_back(devices[g]);
    }
    int k = 0, r = 0;
    b200v_codec_shape(codec, &k, &r);
    s->sym_row = frame_steps * (r ? r : R) * b200v_codec_symbol_bytes(codec);
    return s;
}

extern "C" void b200v_sharded_destroy(b200v_sharded* s) {
    if (!s) return;
    for (b200v_batch* p : s->inst) b200v_batch_destroy(p);
    delete s;
}

extern "C" int b200v_sharded_num_shards(const b200v_sharded* s) { return s ? (int)s->inst.size() : 0; }

extern "C" int b200v_sharded_shard(const b200v_sharded* s, int shard, int* frame0, int* count) {
    if (!s || shard < 0 || shard >= (int)s->inst.size()) return B200V_ERR_ARG;
    if (frame0) *frame0 = s->frame0[(size_t)shard];
    if (count) *count = s->count[(size_t)shard];
    return B200V_OK;
}

extern "C" const char* b200v_sharded_last_error(const b200v_sharded* s) { return s ? s->error.c_str() : ""; }

// init + update + chainback of every shard from host buffers; syms [n_frames][frame_steps*R], data
// [n_frames][(nbits+7)/8], densely packed. Returns the first failing shard's status.
extern "C" int b200v_sharded_decode(b200v_sharded* s, const void* syms, uint8_t* data, unsigned nbits) {
    if (!s || !syms || (!data && nbits)) return B200V_ERR_ARG;
    const size_t orow = (nbits + 7) / 8;
    const size_t n = s->inst.size();
    std::vector<int> rc(n, B200V_OK);
    std::vector<std::string> msg(n);
    std::vector<std::thread> th;
    for (size_t g = 0; g < n; g++) {
        if (!s->inst[g]) continue;
        th.emplace_back([&, g]() {
            b200v_bind_host_to_device(s->device[g]);      // the thread that issues this shard's copies runs next to its GPU
            rc[g] = b200v_batch_decode(s->inst[g], (const char*)syms + (size_t)s->frame0[g] * (size_t)s->sym_row,
                                       data + (size_t)s->frame0[g] * orow, nbits);
            if (rc[g] != B200V_OK) msg[g] = b200v_last_error();      // thread-local in the library
        });
    }
    for (std::thread& t : th) t.join();
    for (size_t g = 0; g < n; g++)
        if (rc[g] != B200V_OK) { s->error = "shard " + std::to_string(g) + ": " + msg[g]; return rc[g]; }
    // pageable buffers work, but the driver stages every chunk copy synchronously: no copy / compute overlap, and the
    // shards' copies serialise in the driver. Not an error -- the note is what b200v_sharded_last_error() returns.
    s->error.clear();
    if (!b200v_host_is_pinned(syms) || (data && !b200v_host_is_pinned(data)))
        s->error = "note: pageable host buffers -- copies were staged synchronously; allocate them with "
                   "b200v_host_alloc_pinned (or cudaHostRegister them) for the pipelined rate";
    return B200V_OK;
}

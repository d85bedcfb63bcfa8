// cassmantle_b200 -- vocabulary lookup on the device (include/...: cm_vocab_*, cm_score_words_host).
//
// the step immediately before the scoring kernel.  In the reference it is
// `inputs.lower()` (src/backend.py:304) followed by gensim's `key_to_index[word]` dict lookup
// inside `wv.similarity` (src/backend.py:307), a KeyError meaning "unknown word".  Here the
// dictionary is an open-addressing hash table in HBM:
//
//   slot   = { uint32 tag (high hash bits, 0 = empty), int32 row }           8 bytes, load <= 0.5
//   keys   = the UTF-8 bytes of every key, concatenated, + int64 offsets     compared on a tag hit
//   hash   = 64-bit FNV-1a over the bytes after ASCII lower-casing
//
// Built once on the host in key order ("first occurrence wins", like a dict built from
// index_to_key), uploaded, read-only afterwards.  One thread looks up one word: hash, linear
// probe, byte compare -- integer work, so results are bit-exact against the Python dict (the
// oracle for this row IS a Python dict).  The same kernel computes the reference's string
// equality `inputs.lower() == answer.lower()` (src/backend.py:305) against the guess's answer.
//
// Lower-casing on the device covers ASCII only (the frontend admits nothing but [a-zA-Z0-9],
// static/script.js:422-428); a caller holding other text lowers it first, as Python would.

#include <cuda_runtime.h>

#include <cstring>
#include <new>
#include <vector>

#include "cm_internal.h"

struct cm_vocab {
    cm_engine *engine = nullptr;
    int64_t n_keys = 0;
    uint64_t n_slots = 0;            // power of two
    uint2 *d_slots = nullptr;        // {tag, row}
    uint8_t *d_keys = nullptr;
    int64_t *d_key_offsets = nullptr;
    // device scratch of cm_score_words_host (grown on demand)
    unsigned char *d_work = nullptr;
    unsigned char *h_work = nullptr;
    size_t work_bytes = 0;
};

namespace {

using cm::fail;

__host__ __device__ __forceinline__ uint8_t ascii_lower(uint8_t c) {
    return (c >= 'A' && c <= 'Z') ? (uint8_t)(c + 32) : c;
}

__host__ __device__ __forceinline__ uint64_t fnv1a_lower(const uint8_t *p, int64_t n) {
    uint64_t h = 0xcbf29ce484222325ull;
    for (int64_t i = 0; i < n; ++i) {
        h ^= (uint64_t)ascii_lower(p[i]);
        h *= 0x100000001b3ull;
    }
    return h;
}

__host__ __device__ __forceinline__ uint32_t tag_of(uint64_t h) { return (uint32_t)(h >> 32) | 1u; }

// rows[i] = row of word i (lower-cased) or -1; optionally exact[i] = (word i == answer of its slot)
__global__ void __launch_bounds__(128)
cm_vocab_lookup_kernel(const uint2 *__restrict__ slots, uint64_t mask,
                       const uint8_t *__restrict__ keys, const int64_t *__restrict__ key_offsets,
                       const uint8_t *__restrict__ bytes, const int32_t *__restrict__ offsets,
                       int n_words, int32_t *__restrict__ rows,
                       const uint8_t *__restrict__ answer_bytes, const int32_t *__restrict__ answer_offsets,
                       int n_answers, const int32_t *__restrict__ answer_slot, uint8_t *__restrict__ exact) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_words) return;
    const int32_t lo = offsets[i];
    const int32_t len = offsets[i + 1] - lo;
    const uint8_t *w = bytes + lo;
    const uint64_t h = fnv1a_lower(w, len);
    const uint32_t tag = tag_of(h);
    int32_t row = -1;
    for (uint64_t s = h & mask;; s = (s + 1) & mask) {
        const uint2 e = slots[s];
        if (e.x == 0u) break;                                   // empty slot: the word is unknown
        if (e.x != tag) continue;
        const int64_t ko = key_offsets[e.y];
        if (key_offsets[e.y + 1] - ko != (int64_t)len) continue;
        bool same = true;
        for (int32_t k = 0; k < len && same; ++k) same = keys[ko + k] == ascii_lower(w[k]);
        if (same) { row = (int32_t)e.y; break; }
    }
    rows[i] = row;
    if (exact != nullptr) {
        const int32_t sl = answer_slot[i];
        bool eq = false;
        if (sl >= 0 && sl < n_answers) {
            const int32_t ao = answer_offsets[sl];
            eq = (answer_offsets[sl + 1] - ao) == len;
            for (int32_t k = 0; k < len && eq; ++k) eq = ascii_lower(answer_bytes[ao + k]) == ascii_lower(w[k]);
        }
        exact[i] = eq ? 1 : 0;
    }
}

int ensure_work(cm_vocab *v, size_t bytes) {
    if (bytes <= v->work_bytes) return CM_OK;
    size_t want = v->work_bytes ? v->work_bytes : (size_t)1 << 20;
    while (want < bytes) want *= 2;
    if (v->d_work) cudaFree(v->d_work);
    if (v->h_work) cudaFreeHost(v->h_work);
    v->d_work = nullptr; v->h_work = nullptr; v->work_bytes = 0;
    if (cudaMalloc(&v->d_work, want) != cudaSuccess || cudaHostAlloc(&v->h_work, want, cudaHostAllocDefault) != cudaSuccess) {
        cudaGetLastError();
        return fail(CM_ERR_NOMEM, "vocabulary work buffers (%zu bytes) failed", want);
    }
    v->work_bytes = want;
    return CM_OK;
}

int launch_lookup(cm_vocab *v, const uint8_t *d_bytes, const int32_t *d_offsets, int n_words,
                  int32_t *d_rows, const uint8_t *d_answer_bytes, const int32_t *d_answer_offsets,
                  int n_answers, const int32_t *d_answer_slot, uint8_t *d_exact, cudaStream_t st) {
    if (n_words == 0) return CM_OK;
    cm_vocab_lookup_kernel<<<(n_words + 127) / 128, 128, 0, st>>>(
        v->d_slots, v->n_slots - 1, v->d_keys, v->d_key_offsets, d_bytes, d_offsets, n_words, d_rows,
        d_answer_bytes, d_answer_offsets, n_answers, d_answer_slot, d_exact);
    cm::count_launch();
    CM_CUDA(cudaGetLastError());
    return CM_OK;
}

}  // namespace

extern "C" {

int cm_vocab_destroy(cm_vocab *v) {
    if (v == nullptr) return CM_OK;
    {
        cm::DeviceGuard guard(v->engine->device);
        if (v->d_slots) cudaFree(v->d_slots);
        if (v->d_keys) cudaFree(v->d_keys);
        if (v->d_key_offsets) cudaFree(v->d_key_offsets);
        if (v->d_work) cudaFree(v->d_work);
        if (v->h_work) cudaFreeHost(v->h_work);
    }
    delete v;
    return CM_OK;
}

int cm_vocab_create(cm_vocab **out, cm_engine *e, const char *keys, const int64_t *key_offsets,
                    int64_t n_keys) {
    if (out == nullptr) return fail(CM_ERR_INVALID, "out is NULL");
    *out = nullptr;
    if (int rc = cm::check_engine(e)) return rc;
    if (keys == nullptr || key_offsets == nullptr) return fail(CM_ERR_INVALID, "keys or key_offsets is NULL");
    if (n_keys < 1 || n_keys > 0x7fffffffLL)
        return fail(CM_ERR_INVALID, "n_keys %lld outside [1, 2^31)", (long long)n_keys);
    for (int64_t i = 0; i < n_keys; ++i)
        if (key_offsets[i + 1] < key_offsets[i])
            return fail(CM_ERR_INVALID, "key_offsets is not non-decreasing at %lld", (long long)i);
    uint64_t n_slots = 16;
    while (n_slots < (uint64_t)n_keys * 2) n_slots <<= 1;
    std::vector<uint2> slots;
    try {
        slots.assign(n_slots, make_uint2(0u, 0u));
    } catch (const std::bad_alloc &) {
        return fail(CM_ERR_NOMEM, "host table of %llu slots failed", (unsigned long long)n_slots);
    }
    const uint64_t mask = n_slots - 1;
    const uint8_t *kb = reinterpret_cast<const uint8_t *>(keys);
    for (int64_t i = 0; i < n_keys; ++i) {           // key order: the first occurrence of a key wins
        const uint8_t *k = kb + key_offsets[i];
        const int64_t len = key_offsets[i + 1] - key_offsets[i];
        // keys are stored and compared verbatim; the QUERY is what gets lower-cased.  Hash the
        // key with the same lower-casing function so a lower-case key lands where its query looks.
        const uint64_t h = fnv1a_lower(k, len);
        const uint32_t tag = tag_of(h);
        bool dup = false;
        uint64_t s = h & mask;
        for (;; s = (s + 1) & mask) {
            if (slots[s].x == 0u) break;
            if (slots[s].x != tag) continue;
            const int64_t o = key_offsets[slots[s].y];
            if (key_offsets[slots[s].y + 1] - o == len && memcmp(kb + o, k, (size_t)len) == 0) { dup = true; break; }
        }
        if (!dup) slots[s] = make_uint2(tag, (uint32_t)i);
    }
    cm_vocab *v = new (std::nothrow) cm_vocab();
    if (v == nullptr) return fail(CM_ERR_NOMEM, "vocabulary allocation failed");
    v->engine = e; v->n_keys = n_keys; v->n_slots = n_slots;
    const size_t key_bytes = (size_t)key_offsets[n_keys] > 0 ? (size_t)key_offsets[n_keys] : 1;
    cm::DeviceGuard guard(e->device);
    cudaError_t err = cudaMalloc(&v->d_slots, n_slots * sizeof(uint2));
    if (err == cudaSuccess) err = cudaMalloc(&v->d_keys, key_bytes);
    if (err == cudaSuccess) err = cudaMalloc(&v->d_key_offsets, (size_t)(n_keys + 1) * sizeof(int64_t));
    if (err == cudaSuccess) err = cudaMemcpy(v->d_slots, slots.data(), n_slots * sizeof(uint2), cudaMemcpyHostToDevice);
    if (err == cudaSuccess && key_offsets[n_keys] > 0)
        err = cudaMemcpy(v->d_keys, keys + key_offsets[0], (size_t)(key_offsets[n_keys] - key_offsets[0]), cudaMemcpyHostToDevice);
    if (err == cudaSuccess) {
        std::vector<int64_t> rel((size_t)n_keys + 1);
        for (int64_t i = 0; i <= n_keys; ++i) rel[(size_t)i] = key_offsets[i] - key_offsets[0];
        err = cudaMemcpy(v->d_key_offsets, rel.data(), rel.size() * sizeof(int64_t), cudaMemcpyHostToDevice);
    }
    if (err != cudaSuccess) {
        cudaGetLastError();
        const int code = fail(CM_ERR_NOMEM, "vocabulary upload failed: %s", cudaGetErrorString(err));
        cm_vocab_destroy(v);
        return code;
    }
    *out = v;
    return CM_OK;
}

int cm_vocab_size(cm_vocab *v, int64_t *n_keys, int64_t *n_slots) {
    if (v == nullptr) return fail(CM_ERR_INVALID, "vocabulary is NULL");
    if (n_keys) *n_keys = v->n_keys;
    if (n_slots) *n_slots = (int64_t)v->n_slots;
    return CM_OK;
}

int cm_vocab_lookup_device(cm_vocab *v, const uint8_t *d_bytes, const int32_t *d_offsets, int n_words,
                           int32_t *d_rows, void *stream) {
    if (v == nullptr) return fail(CM_ERR_INVALID, "vocabulary is NULL");
    if (n_words < 0) return fail(CM_ERR_INVALID, "n_words %d is negative", n_words);
    if (n_words > 0 && (!d_bytes || !d_offsets || !d_rows)) return fail(CM_ERR_INVALID, "a required pointer is NULL");
    cm::DeviceGuard guard(v->engine->device);
    return launch_lookup(v, d_bytes, d_offsets, n_words, d_rows, nullptr, nullptr, 0, nullptr, nullptr,
                         (cudaStream_t)stream);
}

int cm_vocab_lookup_host(cm_vocab *v, const uint8_t *bytes, const int32_t *offsets, int n_words, int32_t *rows) {
    if (v == nullptr) return fail(CM_ERR_INVALID, "vocabulary is NULL");
    if (n_words < 0) return fail(CM_ERR_INVALID, "n_words %d is negative", n_words);
    if (n_words == 0) return CM_OK;
    if (!bytes || !offsets || !rows) return fail(CM_ERR_INVALID, "a required pointer is NULL");
    if (offsets[0] != 0) return fail(CM_ERR_INVALID, "offsets[0] must be 0");
    cm_engine *e = v->engine;
    std::lock_guard<std::mutex> lock(e->mu);
    cm::DeviceGuard guard(e->device);
    const size_t n_bytes = (size_t)offsets[n_words];
    const size_t off_offsets = cm::align_up(n_bytes + 1, 16);
    const size_t off_rows = off_offsets + cm::align_up((size_t)(n_words + 1) * 4, 16);
    const size_t total = off_rows + (size_t)n_words * 4;
    if (int rc = ensure_work(v, total)) return rc;
    memcpy(v->h_work, bytes, n_bytes);
    memcpy(v->h_work + off_offsets, offsets, (size_t)(n_words + 1) * 4);
    CM_CUDA(cudaMemcpyAsync(v->d_work, v->h_work, off_rows, cudaMemcpyHostToDevice, e->stream));
    if (int rc = launch_lookup(v, v->d_work, (const int32_t *)(v->d_work + off_offsets), n_words,
                               (int32_t *)(v->d_work + off_rows), nullptr, nullptr, 0, nullptr, nullptr, e->stream))
        return rc;
    CM_CUDA(cudaMemcpyAsync(rows, v->d_work + off_rows, (size_t)n_words * 4, cudaMemcpyDeviceToHost, e->stream));
    CM_CUDA(cudaStreamSynchronize(e->stream));
    return CM_OK;
}

// Backend.compute_scores (src/backend.py:312-317) for one-word guesses with the strings going
// to the device as they are: lower(), equality, key lookup, similarity, floor -- three launches
// (answers' lookup + unit vectors, guesses' lookup, scoring) on the engine's stream.
int cm_score_words_host(cm_engine *e, cm_vocab *v, const uint8_t *guess_bytes, const int32_t *guess_offsets,
                        int n_guesses, const uint8_t *answer_bytes, const int32_t *answer_offsets,
                        int n_answers, const int32_t *answer_slot, float min_score, float *scores,
                        int32_t *centi) {
    if (v == nullptr) return fail(CM_ERR_INVALID, "vocabulary is NULL");
    if (v->engine != e) return fail(CM_ERR_INVALID, "vocabulary belongs to another engine");
    if (n_guesses < 0) return fail(CM_ERR_INVALID, "n_guesses %d is negative", n_guesses);
    if (n_answers < 1 || n_answers > CM_MAX_ANSWERS)
        return fail(CM_ERR_INVALID, "n_answers %d outside [1, %d]", n_answers, CM_MAX_ANSWERS);
    if (n_guesses == 0) return CM_OK;
    if (!guess_bytes || !guess_offsets || !answer_bytes || !answer_offsets || !answer_slot || !scores)
        return fail(CM_ERR_INVALID, "a required pointer is NULL");
    if (guess_offsets[0] != 0 || answer_offsets[0] != 0) return fail(CM_ERR_INVALID, "offsets must start at 0");
    std::lock_guard<std::mutex> lock(e->mu);
    cm::DeviceGuard guard(e->device);
    const size_t B = (size_t)n_guesses, A = (size_t)n_answers;
    const size_t gb = (size_t)guess_offsets[B], ab = (size_t)answer_offsets[A];
    // packed input block (one DMA): guess bytes | guess offsets | answer bytes | answer offsets | slots
    size_t o = 0;
    const size_t o_gb = o;      o = cm::align_up(o + gb + 1, 16);
    const size_t o_go = o;      o = cm::align_up(o + (B + 1) * 4, 16);
    const size_t o_ab = o;      o = cm::align_up(o + ab + 1, 16);
    const size_t o_ao = o;      o = cm::align_up(o + (A + 1) * 4, 16);
    const size_t o_sl = o;      o = cm::align_up(o + B * 4, 16);
    const size_t in_bytes = o;
    // device-only: answer rows | guess rows | guess len | exact ; then outputs: scores | centi
    const size_t o_ar = o;      o = cm::align_up(o + A * 4, 16);
    const size_t o_gr = o;      o = cm::align_up(o + B * 4, 16);
    const size_t o_gl = o;      o = cm::align_up(o + B * 4, 16);
    const size_t o_ex = o;      o = cm::align_up(o + B, 16);
    const size_t o_sc = o;      o = cm::align_up(o + B * 4, 16);
    const size_t o_ce = o;      o = cm::align_up(o + B * 4, 16);
    if (int rc = ensure_work(v, o)) return rc;
    unsigned char *h = v->h_work, *d = v->d_work;
    memcpy(h + o_gb, guess_bytes, gb);
    memcpy(h + o_go, guess_offsets, (B + 1) * 4);
    memcpy(h + o_ab, answer_bytes, ab);
    memcpy(h + o_ao, answer_offsets, (A + 1) * 4);
    memcpy(h + o_sl, answer_slot, B * 4);
    int32_t *ones = (int32_t *)(h + o_gl);
    for (size_t i = 0; i < B; ++i) ones[i] = 1;                 // every guess is one word
    cudaStream_t st = e->stream;
    CM_CUDA(cudaMemcpyAsync(d, h, in_bytes, cudaMemcpyHostToDevice, st));
    CM_CUDA(cudaMemcpyAsync(d + o_gl, h + o_gl, B * 4, cudaMemcpyHostToDevice, st));
    // 1. answers: rows, then unit vectors (cm_set_answers_device)
    if (int rc = launch_lookup(v, d + o_ab, (const int32_t *)(d + o_ao), n_answers, (int32_t *)(d + o_ar),
                               nullptr, nullptr, 0, nullptr, nullptr, st))
        return rc;
    if (int rc = cm_set_answers_device(e, (const int32_t *)(d + o_ar), n_answers, st)) return rc;
    // 2. guesses: rows + string equality with their answer
    if (int rc = launch_lookup(v, d + o_gb, (const int32_t *)(d + o_go), n_guesses, (int32_t *)(d + o_gr),
                               d + o_ab, (const int32_t *)(d + o_ao), n_answers, (const int32_t *)(d + o_sl),
                               d + o_ex, st))
        return rc;
    // 3. score
    if (int rc = cm::enqueue_score(e, (const int32_t *)(d + o_gr), (const int32_t *)(d + o_gl), n_guesses, 1,
                                   (const int32_t *)(d + o_sl), d + o_ex, min_score, (float *)(d + o_sc),
                                   centi ? (int32_t *)(d + o_ce) : nullptr, 0u, st))
        return rc;
    CM_CUDA(cudaMemcpyAsync(h + o_sc, d + o_sc, (centi ? o_ce + B * 4 : o_sc + B * 4) - o_sc, cudaMemcpyDeviceToHost, st));
    CM_CUDA(cudaStreamSynchronize(st));
    memcpy(scores, h + o_sc, B * 4);
    if (centi) memcpy(centi, h + o_ce, B * 4);
    return CM_OK;
}

}  // extern "C"

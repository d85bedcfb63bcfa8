/*
 * cassmantle_b200 -- C ABI of the B200-native CassMantle guess scorer.
 *
 * This header is the drop-in boundary for ONE path of the reference
 * (SnowCheetos/CassMantle): Backend.compute_score / Backend.compute_scores,
 * src/backend.py:303-317, whose arithmetic is gensim's KeyedVectors.similarity
 * (call site src/backend.py:307) over the table loaded at src/backend.py:45.
 * The reference is Python; the binding a maintainer adds is the ctypes stub in
 * INTEGRATION.md (shipped here as cassmantle_b200/_lib.py).
 *
 * Conventions
 *   - plain C types only; no exceptions cross the boundary; every call returns
 *     CM_OK (0) or a negative cm_status, and cm_last_error() gives the text.
 *   - "d_" arguments are device pointers on the engine's GPU, others are host pointers.
 *   - a guess is N word slots: guess_idx[b*N + k] is a table row, -1 means "not in the
 *     vocabulary" (KeyError in the reference).  guess_len[b] in [0, N] says how many
 *     slots are used; the rest is padding.  N = 1 is the only shape the reference
 *     itself produces; N > 1 is the mean-pool extension (gensim n_similarity).
 *   - answers are registered once per round (cm_set_answers) and referred to by slot.
 *   - `stream` is a cudaStream_t passed as void*, used as given: NULL is CUDA's default
 *     stream, exactly as in the runtime API.  cm_engine_stream() returns the engine's own
 *     non-blocking stream, which the host-buffer entry point uses.
 */
#ifndef CASSMANTLE_B200_H
#define CASSMANTLE_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct cm_engine cm_engine;
typedef struct cm_pipe cm_pipe;
typedef struct cm_gather cm_gather;
typedef struct cm_vocab cm_vocab;
typedef struct cm_blur cm_blur;

typedef enum cm_status {
    CM_OK = 0,
    CM_ERR_INVALID = -1,   /* bad argument (null pointer, size out of range, ...)  */
    CM_ERR_CUDA = -2,      /* a CUDA runtime call failed; see cm_last_error()      */
    CM_ERR_NOMEM = -3,     /* host or device allocation failed                     */
    CM_ERR_STATE = -4      /* call order violated (e.g. scoring before answers)    */
} cm_status;

#define CM_MAX_DIM 1024        /* vector_size limit; must be a multiple of 4            */
#define CM_MAX_ANSWERS 4096    /* answer slots per engine                               */
#define CM_MAX_SEQ 4096        /* word slots per guess                                  */

/* cm_score_batch_device flags.
 * CM_SCORE_OVERLAP_PREVIOUS: the launch may start gathering table rows while the kernel
 * launched just before it on the same stream is still finishing (programmatic dependent
 * launch); it still waits for that kernel before reading the answers or writing results.
 * Pass it only when the guess buffers of this call (d_guess_idx, d_guess_len,
 * d_answer_slot, d_exact) were NOT written by that immediately preceding kernel -- e.g. a
 * stream of independent batches, each filled by a copy or by earlier work.              */
#define CM_SCORE_OVERLAP_PREVIOUS 0x1u

/* Text of the last error raised on the calling thread ("" if none). */
const char *cm_last_error(void);

/* Library version, "major.minor.patch". */
const char *cm_version(void);

/* ---- lifecycle ---------------------------------------------------------------------- *
 * Replaces `self.wv = KeyedVectors.load("data/word2vec.wordvectors", mmap='r')`,
 * src/backend.py:45: the n_rows x dim fp32 matrix becomes resident in the GPU's HBM.      */
int cm_engine_create(cm_engine **out, int device, int64_t n_rows, int dim);
int cm_engine_destroy(cm_engine *e);   /* destroy its pipes, gathers and vocabularies first */

/* Copy rows [row0, row0+n) of the host matrix (row-major fp32, e.g. a slice of gensim's
 * mmap'd `wv.vectors`) into the resident table.  Blocks until the copy is done.          */
int cm_engine_upload_rows(cm_engine *e, const float *rows, int64_t row0, int64_t n);

/* Fill the resident table with the seeded synthetic stand-in (cassmantle_b200/synth.py
 * produces the same bits on the host).  No reference counterpart: the real weights
 * (download_model.py:9) cannot be fetched offline.                                       */
int cm_engine_fill_synthetic(cm_engine *e, uint32_t seed);

/* Device pointer / geometry of the resident table (for tests and stream plumbing).       */
int cm_engine_table(cm_engine *e, void **d_table, int64_t *n_rows, int *dim);

/* Copy table rows back to the host (tests: compare the synthetic fill with synth.py).    */
int cm_engine_download_rows(cm_engine *e, float *rows, int64_t row0, int64_t n);

/* Measurement aid: enqueue a pure read of the first 512 MB of the resident table on `stream`, which
 * leaves the 126 MB L2 full of clean lines (benchmarks call it before a timed region).           */
int cm_engine_flush_l2(cm_engine *e, void *stream);

/* The engine's own stream as a cudaStream_t, and a blocking wait on it.                  */
int cm_engine_stream(cm_engine *e, void **stream);
int cm_engine_synchronize(cm_engine *e);

/* ---- answers (src/server.py:65-72 builds them: prompt['tokens'][int(m)]) ------------- *
 * Registers the round's hidden words as table rows (-1 = not in the vocabulary) and
 * precomputes unitvec(row) for each, exactly once, the way gensim's similarity() would on
 * every call.  n_answers in [1, CM_MAX_ANSWERS].  Asynchronous on `stream`.              */
int cm_set_answers(cm_engine *e, const int32_t *answer_rows, int n_answers, void *stream);
int cm_set_answers_device(cm_engine *e, const int32_t *d_answer_rows, int n_answers, void *stream);

/* ---- the hot path --------------------------------------------------------------------- *
 * Batched Backend.compute_score (src/backend.py:303-310) for B guesses of N slots:
 *   exact[b] != 0                      -> 1.0        (string equality, backend.py:305)
 *   any used slot < 0, len 0, or the
 *   answer row < 0                     -> min_score  (KeyError, backend.py:308-309)
 *   otherwise cos = dot(unitvec(mean of rows), unitvec(answer row))   (backend.py:307)
 *   result = cos if cos > min_score else min_score                    (backend.py:310)
 * d_exact may be NULL (no exact matches).  d_centi may be NULL; when given it receives
 * floor(score * 1e4 + 0.5), the browser's (score*100).toFixed(2) in hundredths
 * (static/script.js:300).  Asynchronous on `stream`; inputs and outputs stay in HBM.
 * `flags`: 0, or CM_SCORE_OVERLAP_PREVIOUS (see above).  The call only enqueues a kernel,
 * so a run of such calls can be captured in a CUDA graph and replayed (tested).            */
int cm_score_batch_device(cm_engine *e,
                          const int32_t *d_guess_idx, const int32_t *d_guess_len,
                          int n_guesses, int n_slots,
                          const int32_t *d_answer_slot, const uint8_t *d_exact,
                          float min_score, float *d_scores, int32_t *d_centi,
                          uint32_t flags, void *stream);

/* The same with 3-byte indices (little-endian row number, 0xFFFFFF = not in the vocabulary): the
 * packed form that the pipelined host path ships across PCIe (cm_pipe_set_idx24).  Tables of at most
 * 16,777,215 rows; the word2vec table (download_model.py:9) has 3,000,000.                          */
int cm_score_batch_device_idx24(cm_engine *e,
                                const uint8_t *d_guess_idx24, const int32_t *d_guess_len,
                                int n_guesses, int n_slots,
                                const int32_t *d_answer_slot, const uint8_t *d_exact,
                                float min_score, float *d_scores, int32_t *d_centi,
                                uint32_t flags, void *stream);

/* Same call with HOST buffers: stages the inputs to the GPU, scores, copies the results
 * back and returns when `scores` (and `centi`, if not NULL) are valid.  This is what the
 * Python boundary (cassmantle_b200/backend.py) calls per micro-batch.  Page-locked
 * buffers (cudaHostAlloc / cudaHostRegister, e.g. torch pin_memory) are read and written
 * by the kernel directly; pageable ones go through one pinned staging block.  `scores`
 * and `centi` are overwritten while the call runs (also on error).                         */
int cm_score_batch_host(cm_engine *e,
                        const int32_t *guess_idx, const int32_t *guess_len,
                        int n_guesses, int n_slots,
                        const int32_t *answer_slot, const uint8_t *exact,
                        float min_score, float *scores, int32_t *centi);

/* Resident kernel for micro-batches.  No reference counterpart: the reference scores a request's
 * two words inline (src/backend.py:312-317); this serves BASELINE configs[1], many small
 * concurrent requests coalesced by the micro-batcher, whose metric is latency.  One blocking
 * cm_score_batch_host call for a few hundred guesses spends most of its time launching a 4 us
 * kernel and synchronising with it.  After cm_engine_arm_resident the engine keeps a small
 * kernel resident (one 128-thread CTA per SM, idle at low duty) that polls a doorbell in
 * page-locked memory: cm_score_batch_host then packs the batch into the mailbox, rings, and
 * watches the results arrive -- two PCIe crossings, no launch.  Used for calls with
 * n_guesses <= 1024 and n_guesses * n_slots <= 4096 (where it is measured to win); larger ones
 * take the launch path as before; results are bit-identical either way.
 * The kernel leaves by itself after `idle_ms` (0 = 20 ms) without a request and is launched
 * again by the next small call, so anything in the process that synchronises the whole device
 * (cudaFree, cudaDeviceSynchronize) waits at most that long for it.  cm_engine_upload_rows and
 * cm_engine_fill_synthetic stop it first.  cm_engine_disarm_resident stops it and returns
 * cm_score_batch_host to the launch path; cm_engine_destroy does so too.                       */
int cm_engine_arm_resident(cm_engine *e, int idle_ms);
int cm_engine_disarm_resident(cm_engine *e);
int cm_engine_resident_trace(cm_engine *e, double *out7);   /* developer aid, see cm_scorer.cu */

/* ---- pipelined host path --------------------------------------------------------------- *
 * No reference counterpart: the reference scores one request at a time inside its event
 * loop (src/backend.py:312-317).  This is the device half of the request micro-batcher that
 * north_star adds to src/server.py.  A pipe is a ring of `depth` slots; each slot owns a
 * page-locked host block that the caller fills IN PLACE.  A step is either ONE kernel that
 * reads the block and writes the scores into it across PCIe (always for inputs up to 64 KiB),
 * or H2D copy, kernel, D2H copy on the slot's stream (always above 8192 guesses or 640 KiB);
 * in between the pipe times both routes as it runs and keeps the faster one (environment
 * CM_PIPE_ROUTE=inplace|copy pins it).  Results do not depend on the route.  The slot's
 * contents must not be touched between submit and wait.
 *
 *   cm_pipe_buffers(p, s, B, N, &idx, &len, &slot, &exact, &scores, &centi)   fill these
 *   cm_pipe_submit(p, s, min_score, want_centi)                                asynchronous
 *   ... fill and submit other slots ...
 *   cm_pipe_wait(p, s)                                   scores[0..B) (and centi) are valid
 *
 * Every entry of exact[] must be written (0 or 1).  Answers must be registered and complete
 * (cm_set_answers + cm_engine_synchronize, or the same stream order) before a submit, and
 * must not be changed while any slot is in flight.  Not thread-safe per pipe.              */
int cm_pipe_create(cm_pipe **out, cm_engine *e, int depth, int max_guesses, int max_slots);
int cm_pipe_destroy(cm_pipe *p);
int cm_pipe_depth(cm_pipe *p);
int cm_pipe_buffers(cm_pipe *p, int slot, int n_guesses, int n_slots,
                    int32_t **guess_idx, int32_t **guess_len, int32_t **answer_slot,
                    uint8_t **exact, float **scores, int32_t **centi);
/* Packed indices: after cm_pipe_set_idx24(p, 1) the guess_idx block of every slot holds 3 bytes per
 * index, little-endian, 0xFFFFFF = not in the vocabulary (cm_pipe_buffers still returns it through
 * the int32_t** argument: treat it as uint8_t[n_guesses * n_slots * 3]).  Same scores; a quarter fewer
 * input bytes per guess across PCIe, which is what bounds several GPUs behind one host.  Tables of at
 * most 16,777,215 rows.  Call cm_pipe_buffers again afterwards.                                    */
int cm_pipe_set_idx24(cm_pipe *p, int on);
int cm_pipe_submit(cm_pipe *p, int slot, float min_score, int want_centi);
int cm_pipe_wait(cm_pipe *p, int slot);
/* cm_pipe_wait(wait_slot) (skipped when wait_slot < 0) followed by cm_pipe_submit(submit_slot). */
int cm_pipe_cycle(cm_pipe *p, int wait_slot, int submit_slot, float min_score, int want_centi);
/* The same for `count` consecutive slots (modulo the depth) in one call: for i < count,
 * cm_pipe_wait((wait_first + i) % depth) then cm_pipe_submit((submit_first + i) % depth).
 * wait_first < 0: no waits; submit_first < 0: no submits.  count <= depth.                  */
int cm_pipe_cycle_many(cm_pipe *p, int wait_first, int submit_first, int count, float min_score,
                       int want_centi);
/* Measurement aid: `steps` steps of the copy-engine route with the kernel left out (H2D copy of the
 * slot's inputs, D2H copy of its scores, `in_flight` slots busy at a time); *seconds receives the wall
 * time.  What the host and the PCIe path alone sustain -- the ceiling of the pipelined figure.      */
int cm_pipe_probe_copies(cm_pipe *p, int steps, int in_flight, double *seconds);

/* ---- vocabulary lookup on the device -------------------------------- *
 * Replaces `inputs.lower()` (src/backend.py:304) + gensim's `key_to_index[word]` dict lookup
 * inside `wv.similarity` (src/backend.py:307): an open-addressing hash table in HBM, built
 * once from the keys in index_to_key order (key i -> row i; the first occurrence of a
 * duplicate key wins, as in a dict).  Keys are UTF-8 bytes, `keys + key_offsets[i]` ..
 * `keys + key_offsets[i+1]`.  A query word is lower-cased (ASCII A-Z only; lower other text
 * first, as Python would) and looked up: its row, or -1 where the reference raises KeyError.
 * Integer work: results are bit-exact against the Python dict.                             */
int cm_vocab_create(cm_vocab **out, cm_engine *e, const char *keys, const int64_t *key_offsets,
                    int64_t n_keys);
int cm_vocab_destroy(cm_vocab *v);
int cm_vocab_size(cm_vocab *v, int64_t *n_keys, int64_t *n_slots);
/* words i = bytes[offsets[i] .. offsets[i+1]), i < n_words; rows[i] = table row or -1.       */
int cm_vocab_lookup_device(cm_vocab *v, const uint8_t *d_bytes, const int32_t *d_offsets,
                           int n_words, int32_t *d_rows, void *stream);
int cm_vocab_lookup_host(cm_vocab *v, const uint8_t *bytes, const int32_t *offsets,
                         int n_words, int32_t *rows);
/* Backend.compute_scores (src/backend.py:312-317) for one-word guesses with the STRINGS
 * handed to the device: lower(), string equality with the answer (-> 1.0), key lookup of
 * guesses and answers (unknown -> min_score), similarity, floor.  guess i is
 * guess_bytes[guess_offsets[i] .. guess_offsets[i+1]) and is scored against answer
 * answer_slot[i] of the n_answers answer strings.  Blocking; re-registers the engine's
 * answers (as cm_set_answers would).                                                        */
int cm_score_words_host(cm_engine *e, cm_vocab *v,
                        const uint8_t *guess_bytes, const int32_t *guess_offsets, int n_guesses,
                        const uint8_t *answer_bytes, const int32_t *answer_offsets, int n_answers,
                        const int32_t *answer_slot, float min_score, float *scores, int32_t *centi);

/* ---- most_similar ------------------------------------------------- *
 * Backend.most_similar, src/backend.py:297-301 = gensim KeyedVectors.most_similar(
 * positive=[word], topn): cosine of the word's row against EVERY row (one pass over the
 * table), then the k best.  `row` is the word's table row (the caller raises for unknown
 * words, as the reference does); rows_out/sims_out receive the k best rows by descending
 * cosine INCLUDING `row` itself -- gensim asks for topn+1 and drops the query, so pass
 * k = topn + 1 and drop it too.  k <= 1024.  Blocking; the first call also computes the row
 * norms (gensim's fill_norms).  NaN cosines rank last.                                       */
int cm_most_similar(cm_engine *e, int32_t row, int k, int32_t *rows_out, float *sims_out, int *n_out);

/* The same for n_queries words in ONE pass over the table (an extension for the micro-batcher: the
 * reference calls most_similar once per word).  rows[q] is query q's table row; its k best land in
 * rows_out / sims_out [q * k .. q * k + n_out[q]), exactly what cm_most_similar(rows[q], k) returns --
 * the same rows and the same fp32 cosines (entries tied at the k-th place may resolve differently).
 * Up to 256 queries share a pass: the table is multiplied with the query block on the tensor cores in
 * tf32 only to FILTER candidates (threshold from a sampled pass, widened by the tf32 error bound), and
 * the candidates are re-scored with the one-query arithmetic.  Tables narrower than 32 or wider than 320
 * columns or shorter than 256 rows, a single query, or CM_SIMILAR_BATCH=0 in the environment take the
 * one-query path in a loop.  Host pointers,
 * blocking.                                                                                      */
int cm_most_similar_batch(cm_engine *e, const int32_t *rows, int n_queries, int k,
                          int32_t *rows_out, float *sims_out, int *n_out);

/* ---- mask selection, numeric half ------------------------------------- *
 * semantic_distance(word, word_list, model), src/utils.py:73-79, for every word of word_list
 * at once (select_descriptive_words calls it in a loop, src/utils.py:91):
 *   out[i] = || table[rows[i]] - mean_j table[rows[j]] ||_2   j < n_mean with rows[j] >= 0,
 *   out[i] = 0 where rows[i] < 0 (`word in model` is false),           for all i < n.
 * word_list is rows[0 .. n_mean); n_mean == n scores the list against itself, as the reference
 * does; rows beyond n_mean are extra words measured against the same mean.
 * numpy's operation order: fp32 running sum in list order, / count, fp32 difference,
 * sqrt(dot).  Host pointers, blocking; n <= CM_MAX_SEQ.  Once per round in the reference.     */
int cm_semantic_distances(cm_engine *e, const int32_t *rows, int n, int n_mean, float *out);

/* ---- per-session image blur ---------------------------------------- *
 * Backend.mask_image, src/backend.py:322-324: image.filter(ImageFilter.GaussianBlur(blur))
 * with blur = score_to_blur(score) (src/backend.py:319-320), run per request by
 * Server.fetch_masked_image (src/server.py:129-133).  Pillow's GaussianBlur is three passes
 * per axis of an extended box blur in 32-bit integer arithmetic; this restates it exactly,
 * so the output bytes are identical to Pillow's.  The round's image is uploaded once
 * (cm_blur_set_image: interleaved uint8, height x width x channels, channels 1/3/4 =
 * Pillow modes L/RGB/RGBA); cm_blur_batch_host blurs it for n sessions at once, request i
 * with Gaussian radius radii[i] (the caller evaluates score_to_blur exactly as the reference
 * does and passes it as a float, which is what Pillow's C entry point receives), and
 * returns n images back to back in `out`.  Blocking.                                       */
int cm_blur_create(cm_blur **out, int device, int max_height, int max_width, int channels,
                   int max_batch);
int cm_blur_destroy(cm_blur *b);
int cm_blur_set_image(cm_blur *b, const uint8_t *pixels, int height, int width);
int cm_blur_batch_host(cm_blur *b, const float *radii, int n, uint8_t *out);
/* The same, without the last host copy: *out points at the n images in the blur's own
 * page-locked result block, valid until the next call on `b` (a consumer that encodes or sends
 * the pixels right away -- src/server.py:131-133 saves them as JPEG -- needs no more).      */
int cm_blur_batch_pinned(cm_blur *b, const float *radii, int n, const uint8_t **out);

/* The blurred images as JPEG FILES, which is what the reference sends: GET /fetch/contents saves
 * the blurred PIL image with Pillow's defaults, main.py:100-103 `image.save(img_io, 'JPEG')` (the
 * same call in src/utils.py:12-16 encode_image), and base64-encodes the bytes (main.py:106).
 * Blur (as above) and libjpeg's baseline encoder restated on the device -- RGB->YCbCr, 4:2:0
 * down-sampling, integer 8x8 DCT, quantisation, standard Huffman tables, JFIF header -- so file i
 * equals, byte for byte, what Pillow writes for request i.  RGB objects only (channels == 3;
 * Pillow refuses RGBA as JPEG, the reference's image is RGB).  quality: 1..100 as in Pillow's
 * `quality=`, 0 = Pillow's default 75 (what the reference uses).  File i is
 * out[offsets[i] .. offsets[i + 1]); offsets has n + 1 entries.  Blocking.
 * cm_blur_batch_jpeg copies into the caller's buffer and returns CM_ERR_NOMEM, with offsets[]
 * filled in (offsets[n] = bytes needed), when `capacity` is too small or out is NULL;
 * cm_blur_batch_jpeg_pinned returns pointers into the object's own page-locked block, valid
 * until the next call on `b`.  ~10x fewer bytes cross PCIe than with cm_blur_batch_host.      */
int cm_blur_batch_jpeg(cm_blur *b, const float *radii, int n, int quality, uint8_t *out,
                       int64_t capacity, int64_t *offsets);
int cm_blur_batch_jpeg_pinned(cm_blur *b, const float *radii, int n, int quality,
                              const uint8_t **out, const int64_t **offsets);

/* ---- several GPUs: score all-gather over NVLink peer memory ---------------------------- *
 * No reference counterpart (the reference has no collective).  SURVEY.md §8e: one process
 * per GPU, a table replica each, every rank scores its slice of the batch and every rank
 * needs all scores.  That is 4 bytes per guess -- pure latency for a library collective -- so
 * the exchange rides inside the scoring launches, one step behind: launch s writes its scores
 * to a staging slot; launch s+1 carries one extra CTA per 4096 scores that copies step s
 * into every rank's receive buffer through peer mappings, publishes a step flag on every
 * peer and waits (bounded, ~2 s) for every rank's flag, while the other CTAs already gather
 * rows for step s+1.  One launch per step; the exchange is never on the critical path.
 * world <= 8, one node.
 *
 *   cm_gather_create(&g, e, rank, world, max_guesses_per_rank, depth)      depth >= 2
 *   cm_gather_export(g, h)  -> exchange the 64-byte handles -> cm_gather_import(g, r, h_r)
 *   per step:  cm_score_batch_gather(e, g, ...same arguments as cm_score_batch_device...)
 *              (d_scores may be NULL; when given it also receives this rank's own scores
 *               when the step is shipped)
 *   consumer:  cm_gather_wait(g, &d_all, stream) ships the latest step if no launch has taken
 *              it along yet and waits for all ranks; work enqueued on `stream` afterwards
 *              sees d_all[r * cm_gather_stride(g) + b] = rank r's score b of that step
 * Use one stream for all calls on a gather.  Every rank must make the same sequence of
 * gathered launches (an empty slice is fine) and consume a step's buffer before the
 * (depth/2)-th following gathered launch (ranks may drift depth/2 steps apart, not more).
 * min_score must not be NaN.
 * CM_GATHER_DEFER_WAIT skips the wait for the peers inside the ship (cm_gather_wait still
 * performs it) and cm_gather_flush ships without waiting: both exist for driving several
 * ranks from ONE thread on ONE GPU (the single-GPU tests), where a kernel must never wait
 * for one that has not been launched.  cm_gather_import_ptr links blocks inside a process. */
#define CM_GATHER_DEFER_WAIT 0x2u
int cm_gather_create(cm_gather **out, cm_engine *e, int rank, int world,
                     int max_guesses_per_rank, int depth);
int cm_gather_destroy(cm_gather *g);
int cm_gather_export(cm_gather *g, void *handle64);
int cm_gather_import(cm_gather *g, int peer_rank, const void *handle64);
int cm_gather_import_ptr(cm_gather *g, int peer_rank, void *d_block);
int cm_gather_block(cm_gather *g, void **d_block, int64_t *bytes);
int cm_gather_stride(cm_gather *g);
int cm_score_batch_gather(cm_engine *e, cm_gather *g,
                          const int32_t *d_guess_idx, const int32_t *d_guess_len,
                          int n_guesses, int n_slots,
                          const int32_t *d_answer_slot, const uint8_t *d_exact,
                          float min_score, float *d_scores, int32_t *d_centi,
                          uint32_t flags, void *stream);
/* k batches in one call: enqueues cm_score_batch_gather (g != NULL) or cm_score_batch_device
 * (g == NULL) for batches[0..n_batches) in order on `stream`, with the same min_score and flags,
 * and stops at the first error.  Device pointers as in those calls (scores may be NULL with a
 * gather).  The reference scores one request's keys in a loop (src/backend.py:314-316); this is
 * the batched form of that loop for a caller that has several device batches ready, so that
 * the host language crosses the boundary once, not k times.                                  */
typedef struct cm_batch {
    const int32_t *guess_idx;
    const int32_t *guess_len;
    const int32_t *answer_slot;
    const uint8_t *exact;        /* may be NULL */
    float *scores;
    int32_t *centi;              /* may be NULL */
    int32_t n_guesses;
    int32_t n_slots;
} cm_batch;
int cm_score_batch_list_device(cm_engine *e, cm_gather *g, const cm_batch *batches, int n_batches,
                               float min_score, uint32_t flags, void *stream);
int cm_gather_result(cm_gather *g, float **d_gathered);
int cm_gather_flush(cm_gather *g, void *stream);
int cm_gather_wait(cm_gather *g, float **d_gathered, void *stream);
int cm_gather_status(cm_gather *g, int *timed_out_rank_plus_1);
int cm_gather_trace(cm_gather *g, unsigned long long *out512);   /* developer aid, see cm_gather.cu */

/* ---- scoring without any allocation ---------------------------------------------------- *
 * No counterpart in the reference.  For callers that must not allocate GPU memory: programs
 * whose device memory belongs entirely to their own allocator or to a memory plan fixed at
 * start-up, and request paths that cannot afford cudaMalloc's device-wide synchronisation.
 * They cannot use a cm_engine (it allocates the table replica and the answers' unit
 * vectors), so this call takes everything from the caller: the fp32 table already on
 * `device` (n_rows x dim, row-major), the answers' table rows as a DEVICE array, the batch
 * and the outputs.  Semantics and arguments are otherwise those of cm_set_answers_device
 * followed by cm_score_batch_device with flags 0, and results are identical.
 * The unit vectors of the answers live in one fixed array inside the library
 * (CM_NOALLOC_MAX_ANSWERS x CM_MAX_DIM floats = 256 KiB, resident on every GPU that loads the
 * library until the process exits); n_answers above CM_NOALLOC_MAX_ANSWERS or dim above
 * CM_MAX_DIM is rejected with CM_ERR_INVALID.  The call itself performs no cudaMalloc,
 * cudaFree or other allocating routine.  The first call on a device checks the device and
 * looks the array's address up (once per device): make that first call at start-up, before
 * any stream capture -- with CUDA's lazy module loading it can also be the moment the runtime
 * loads this library onto that device, which allocates and synchronises inside the driver.
 * The address is then cached until the process exits: after cudaDeviceReset() on that device
 * it is stale, so do not call this entry point on a device that has been reset.
 * All calls share the array and each rewrites the part it uses, so calls that may overlap on
 * the GPU (different streams, unordered) race -- keep them on one stream or order them.
 * Asynchronous on `stream`.                                                                 */
#define CM_NOALLOC_MAX_ANSWERS 64
int cm_score_batch_noalloc(int device, const float *d_table, int64_t n_rows, int dim,
                           const int32_t *d_answer_rows, int n_answers,
                           const int32_t *d_guess_idx, const int32_t *d_guess_len,
                           int n_guesses, int n_slots,
                           const int32_t *d_answer_slot, const uint8_t *d_exact,
                           float min_score, float *d_scores, int32_t *d_centi, void *stream);

/* Number of kernels this library has launched since it was loaded (bench.py reports it). */
int64_t cm_kernel_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* CASSMANTLE_B200_H */

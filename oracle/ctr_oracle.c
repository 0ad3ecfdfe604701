/* CPU oracle, plain C restatement of the integer/byte parts of the hot path.
 * TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): linked by tests/ and by bench.py's
 * cpu_baseline / --impl reference legs, never by the product.  PARITY UNPINNED: the
 * reference holds no golden vectors; what is restated here from it is cited per function
 * (DC = /root/reference/src/main/java/tensorflow/serving/DCNClient.java,
 *  P/ = /root/reference/src/main/proto/).
 *
 * Build: make -C oracle      ->  oracle/_build/libctr_oracle.so
 */
#include <stdint.h>
#include <stddef.h>
#include <string.h>
#include <math.h>

#ifdef _OPENMP
#include <omp.h>
#endif

/* ---- DC:46-55 partitionList: range = N / partNum; last part takes the remainder. ---- */
void oracle_partition_bounds(int64_t n, int32_t part_num, int64_t *begin, int64_t *end) {
    int64_t range = n / part_num;
    for (int32_t i = 0; i < part_num - 1; ++i) { begin[i] = i * range; end[i] = (i + 1) * range; }
    begin[part_num - 1] = (int64_t)(part_num - 1) * range;
    end[part_num - 1] = n;
}

/* ---- DC:57-74 demo inputs: ids 1..F per candidate, weights 1.0f. ---- */
void oracle_demo_inputs(int64_t n_cand, int32_t field_num, int64_t *ids, float *wts) {
    for (int64_t c = 0; c < n_cand; ++c)
        for (int32_t f = 0; f < field_num; ++f) {
            ids[c * field_num + f] = f + 1;
            wts[c * field_num + f] = 1.0f;
        }
}

/* ---- builder-defined integer hash (splitmix64 finaliser); bit-exact with the CUDA path ---- */
static inline uint64_t mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
uint64_t oracle_mix64(uint64_t z) { return mix64(z); }

void oracle_id_to_row(const int64_t *ids, int64_t n, int64_t num_rows, int32_t hashed, uint64_t *rows) {
    for (int64_t i = 0; i < n; ++i)
        rows[i] = hashed ? (mix64((uint64_t)ids[i]) & (uint64_t)(num_rows - 1)) : (uint64_t)ids[i];
}

static inline float proc_unit(uint64_t key, uint64_t idx) {
    uint64_t u = mix64(key + idx);
    return (float)(u >> 40) * 0x1p-23f - 1.0f;       /* exact in fp32 */
}
static inline uint64_t proc_key(uint64_t seed, uint64_t tag) { return mix64(seed * 0x100000001B3ull + tag); }

void oracle_proc_unit(uint64_t seed, uint64_t tag, const uint64_t *idx, int64_t n, float *out) {
    uint64_t key = proc_key(seed, tag);
    for (int64_t i = 0; i < n; ++i) out[i] = proc_unit(key, idx[i]);
}

/* Materialise a procedural table in host RAM (the CPU baseline gathers from real memory,
 * as a TF-Serving embedding_lookup would). */
void oracle_fill_table(uint64_t seed, float scale_e, float scale_w1, int64_t num_rows, int32_t K,
                       float *emb, float *w1) {
    uint64_t ke = proc_key(seed, 1), kw = proc_key(seed, 2);
#pragma omp parallel for schedule(static)
    for (int64_t r = 0; r < num_rows; ++r) {
        for (int32_t k = 0; k < K; ++k) emb[r * K + k] = proc_unit(ke, (uint64_t)r * K + k) * scale_e;
        w1[r] = proc_unit(kw, (uint64_t)r) * scale_w1;
    }
}

/* Gather + weight-scale + first-order (+FM) in fp32, sequential sums per candidate. */
void oracle_gather_fp32(const int64_t *ids, const float *wts, int64_t n, int32_t F, int32_t K,
                        int64_t num_rows, int32_t hashed, const float *emb, const float *w1,
                        float b0, int32_t with_fm, float *x0, float *wide) {
#pragma omp parallel for schedule(static)
    for (int64_t b = 0; b < n; ++b) {
        float lin = b0, s2 = 0.f, s1[64];
        for (int32_t k = 0; k < K; ++k) s1[k] = 0.f;
        for (int32_t f = 0; f < F; ++f) {
            uint64_t row = hashed ? (mix64((uint64_t)ids[b * F + f]) & (uint64_t)(num_rows - 1))
                                  : (uint64_t)ids[b * F + f];
            float wt = wts[b * F + f];
            lin += w1[row] * wt;
            const float *e = emb + row * K;
            float *x = x0 + ((size_t)b * F + f) * K;
            for (int32_t k = 0; k < K; ++k) { float v = e[k] * wt; x[k] = v; s1[k] += v; s2 += v * v; }
        }
        if (with_fm) { float q = 0.f; for (int32_t k = 0; k < K; ++k) q += s1[k] * s1[k]; lin += 0.5f * (q - s2); }
        wide[b] = lin;
    }
}

/* ---- protobuf wire format, hand-restated from the field numbers the reference uses ----
 * TensorProto: dtype=1 (varint), tensor_shape=2 (msg), float_val=5 (packed fixed32),
 * int64_val=10 (packed varint)            P/tensorflow/core/framework/tensor.proto:14-84 (:48, :64)
 * TensorShapeProto: dim=2 {size=1}        P/tensorflow/core/framework/tensor_shape.proto:13-46
 * DCNClient fills int64_val / float_val (not tensor_content)          DC:100-101, DC:106-107 */
static inline size_t put_varint(uint8_t *p, uint64_t v) {
    size_t n = 0;
    while (v >= 0x80) { p[n++] = (uint8_t)(v | 0x80); v >>= 7; }
    p[n++] = (uint8_t)v;
    return n;
}
static inline size_t get_varint(const uint8_t *p, uint64_t *v) {
    uint64_t r = 0; size_t n = 0; int sh = 0;
    for (;;) { uint8_t b = p[n++]; r |= (uint64_t)(b & 0x7F) << sh; if (!(b & 0x80)) break; sh += 7; }
    *v = r; return n;
}

/* packed varint payload of an int64 array; returns bytes written (out needs 10*n bytes) */
int64_t oracle_encode_int64_packed(const int64_t *v, int64_t n, uint8_t *out) {
    size_t o = 0;
    for (int64_t i = 0; i < n; ++i) o += put_varint(out + o, (uint64_t)v[i]);
    return (int64_t)o;
}
int64_t oracle_decode_int64_packed(const uint8_t *in, int64_t nbytes, int64_t *v, int64_t cap) {
    size_t o = 0; int64_t i = 0;
    while ((int64_t)o < nbytes && i < cap) { uint64_t x; o += get_varint(in + o, &x); v[i++] = (int64_t)x; }
    return i;
}

static size_t put_tag(uint8_t *p, uint32_t field, uint32_t wt) { return put_varint(p, ((uint64_t)field << 3) | wt); }

/* One TensorProto body with shape [rows, cols]; dtype 9 = DT_INT64 (int64_val), 1 = DT_FLOAT
 * (float_val)                                            P/tensorflow/core/framework/types.proto:18,26 */
static size_t put_tensor(uint8_t *p, int32_t dtype, int64_t rows, int64_t cols,
                         const int64_t *iv, const float *fv, uint8_t *scratch) {
    size_t o = 0, n = (size_t)(rows * cols);
    o += put_tag(p + o, 1, 0); o += put_varint(p + o, (uint64_t)dtype);
    uint8_t shp[64]; size_t s = 0;
    int64_t dims[2] = {rows, cols};
    for (int d = 0; d < 2; ++d) {
        uint8_t dm[16]; size_t ds = 0;
        ds += put_tag(dm + ds, 1, 0); ds += put_varint(dm + ds, (uint64_t)dims[d]);
        s += put_tag(shp + s, 2, 2); s += put_varint(shp + s, ds); memcpy(shp + s, dm, ds); s += ds;
    }
    o += put_tag(p + o, 2, 2); o += put_varint(p + o, s); memcpy(p + o, shp, s); o += s;
    if (dtype == 9) {
        int64_t nb = oracle_encode_int64_packed(iv, (int64_t)n, scratch);
        o += put_tag(p + o, 10, 2); o += put_varint(p + o, (uint64_t)nb); memcpy(p + o, scratch, (size_t)nb); o += (size_t)nb;
    } else {
        o += put_tag(p + o, 5, 2); o += put_varint(p + o, n * 4); memcpy(p + o, fv, n * 4); o += n * 4;
    }
    return o;
}

static size_t put_map_entry(uint8_t *p, uint32_t field, const char *key, const uint8_t *val, size_t vlen) {
    size_t kl = strlen(key), o = 0;
    uint8_t hdr[32]; size_t h = 0;
    h += put_tag(hdr + h, 1, 2); h += put_varint(hdr + h, kl);
    size_t entry = h + kl + 1 + 0;
    uint8_t vl[12]; size_t vls = put_varint(vl, vlen);
    entry += vls + vlen;
    o += put_tag(p + o, field, 2); o += put_varint(p + o, entry);
    memcpy(p + o, hdr, h); o += h; memcpy(p + o, key, kl); o += kl;
    o += put_tag(p + o, 2, 2); memcpy(p + o, vl, vls); o += vls; memcpy(p + o, val, vlen); o += vlen;
    return o;
}

/* PredictRequest{model_spec=1{name=1, signature_name=3}, inputs=2 map<string,TensorProto>}
 * P/tensorflow_serving/apis/predict.proto:12-31, model.proto:9-19; built as DC:91-109 does.
 * out needs about 16*n + 256 bytes; tmp needs 2 * (14*n + 128).  Returns bytes written. */
int64_t oracle_encode_predict_request(const char *model_name, const char *signature,
                                      const int64_t *ids, const float *wts, int64_t rows, int64_t cols,
                                      uint8_t *out, uint8_t *tmp) {
    size_t o = 0, n = (size_t)(rows * cols);
    uint8_t ms[256]; size_t m = 0, l;
    l = strlen(model_name); m += put_tag(ms + m, 1, 2); m += put_varint(ms + m, l); memcpy(ms + m, model_name, l); m += l;
    l = strlen(signature);  m += put_tag(ms + m, 3, 2); m += put_varint(ms + m, l); memcpy(ms + m, signature, l);  m += l;
    o += put_tag(out + o, 1, 2); o += put_varint(out + o, m); memcpy(out + o, ms, m); o += m;
    uint8_t *tb = tmp, *scratch = tmp + 14 * n + 128;
    size_t t = put_tensor(tb, 9, rows, cols, ids, NULL, scratch);
    o += put_map_entry(out + o, 2, "feat_ids", tb, t);
    t = put_tensor(tb, 1, rows, cols, NULL, wts, scratch);
    o += put_map_entry(out + o, 2, "feat_wts", tb, t);
    return (int64_t)o;
}

/* Parse what oracle_encode_predict_request wrote (the server side of DC:112): walks the
 * message generically, fills ids/wts (capacity cap) and the declared shape. Returns number
 * of candidates (rows) or -1 on a malformed message. */
static int64_t parse_tensor(const uint8_t *p, size_t len, int64_t *iv, float *fv, int64_t cap, int64_t dims[2]) {
    size_t o = 0; int64_t count = -1; int nd = 0;
    while (o < len) {
        uint64_t tag, l; o += get_varint(p + o, &tag);
        uint32_t field = (uint32_t)(tag >> 3), wt = (uint32_t)(tag & 7);
        if (wt == 0) { uint64_t v; o += get_varint(p + o, &v); continue; }
        if (wt != 2) return -1;
        o += get_varint(p + o, &l);
        if (field == 2) {                      /* tensor_shape */
            size_t q = 0;
            while (q < l) {
                uint64_t t2, l2; q += get_varint(p + o + q, &t2); q += get_varint(p + o + q, &l2);
                size_t r = 0; uint64_t size = 0;
                while (r < l2) { uint64_t t3; r += get_varint(p + o + q + r, &t3);
                    if ((t3 >> 3) == 1) r += get_varint(p + o + q + r, &size); else return -1; }
                if (nd < 2) dims[nd++] = (int64_t)size;
                q += l2;
            }
        } else if (field == 10 && iv) {
            count = oracle_decode_int64_packed(p + o, (int64_t)l, iv, cap);
        } else if (field == 5 && fv) {
            count = (int64_t)(l / 4); if (count > cap) return -1; memcpy(fv, p + o, l);
        }
        o += l;
    }
    return count;
}

int64_t oracle_decode_predict_request(const uint8_t *in, int64_t nbytes, int64_t *ids, float *wts,
                                      int64_t cap, int64_t shape[2]) {
    size_t o = 0; int64_t nid = -1, nwt = -1; int64_t d1[2] = {0, 0}, d2[2] = {0, 0};
    while ((int64_t)o < nbytes) {
        uint64_t tag, l; o += get_varint(in + o, &tag);
        if ((tag & 7) != 2) return -1;
        o += get_varint(in + o, &l);
        if ((tag >> 3) == 2) {                 /* inputs map entry */
            size_t q = 0; const uint8_t *e = in + o; char key[32] = {0}; const uint8_t *val = NULL; size_t vlen = 0;
            while (q < l) {
                uint64_t t2, l2; q += get_varint(e + q, &t2); q += get_varint(e + q, &l2);
                if ((t2 >> 3) == 1) { size_t c = l2 < 31 ? l2 : 31; memcpy(key, e + q, c); key[c] = 0; }
                else if ((t2 >> 3) == 2) { val = e + q; vlen = l2; }
                q += l2;
            }
            if (val && !strcmp(key, "feat_ids")) nid = parse_tensor(val, vlen, ids, NULL, cap, d1);
            else if (val && !strcmp(key, "feat_wts")) nwt = parse_tensor(val, vlen, NULL, wts, cap, d2);
        }
        o += l;
    }
    if (nid < 0 || nwt < 0 || nid != nwt || d1[0] * d1[1] != nid || d2[0] != d1[0] || d2[1] != d1[1]) return -1;
    shape[0] = d1[0]; shape[1] = d1[1];
    return d1[0];
}

/* PredictResponse{outputs=1 map<string,TensorProto>}: scores as float_val under
 * "prediction_node" (DC:35, DC:162)           P/tensorflow_serving/apis/predict.proto:33-40 */
int64_t oracle_encode_predict_response(const float *scores, int64_t n, uint8_t *out, uint8_t *tmp) {
    size_t t = put_tensor(tmp, 1, n, 1, NULL, scores, NULL);
    return (int64_t)put_map_entry(out, 1, "prediction_node", tmp, t);
}
int64_t oracle_decode_predict_response(const uint8_t *in, int64_t nbytes, float *scores, int64_t cap) {
    size_t o = 0; int64_t dims[2] = {0, 0};
    while ((int64_t)o < nbytes) {
        uint64_t tag, l; o += get_varint(in + o, &tag); o += get_varint(in + o, &l);
        if ((tag >> 3) == 1) {
            size_t q = 0; const uint8_t *e = in + o;
            while (q < l) {
                uint64_t t2, l2; q += get_varint(e + q, &t2); q += get_varint(e + q, &l2);
                if ((t2 >> 3) == 2) return parse_tensor(e + q, l2, NULL, scores, cap, dims);
                q += l2;
            }
        }
        o += l;
    }
    return -1;
}

/* ---- DC:195 Collections.sort on boxed Float: Float.compareTo total order as an int32 key ---- */
static inline int32_t java_key(float f) {
    int32_t b; memcpy(&b, &f, 4);
    if (f != f) b = 0x7FC00000;
    return b < 0 ? (b ^ 0x7FFFFFFF) : b;
}
void oracle_java_sort_keys(const float *x, int64_t n, int32_t *keys) {
    for (int64_t i = 0; i < n; ++i) keys[i] = java_key(x[i]);
}

/* ---- LR closed form (known-answer anchor), fp32 sequential ---- */
void oracle_lr_scores(const int64_t *ids, const float *wts, int64_t n, int32_t F,
                      const float *w1, float b0, float *scores) {
    for (int64_t b = 0; b < n; ++b) {
        float z = b0;
        for (int32_t f = 0; f < F; ++f) z += w1[ids[b * F + f]] * wts[b * F + f];
        scores[b] = (float)(1.0 / (1.0 + exp(-(double)z)));
    }
}

int32_t oracle_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

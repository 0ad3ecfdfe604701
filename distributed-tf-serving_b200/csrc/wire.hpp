// wire.hpp -- TF-Serving Predict wire format, host side.
//
// Lets the engine stand in for the RPC the reference client issues (DCNClient.java:91-115): it takes
// a serialised PredictRequest exactly as DCNClient builds it and returns a serialised
// PredictResponse the client's `getOutputsMap().get("prediction_node").getFloatValList()`
// (DCNClient.java:162) can read.  Hand-written from the field numbers of the reference's vendored
// protos (no protoc on this image):
//   PredictRequest  { model_spec = 1; map<string, TensorProto> inputs = 2; }      tensorflow_serving/apis/predict.proto:12-31
//   PredictResponse { map<string, TensorProto> outputs = 1; model_spec = 2; }     predict.proto:33-40
//   TensorProto     { dtype = 1; tensor_shape = 2; tensor_content = 4; float_val = 5 [packed]; int64_val = 10 [packed]; }
//                                                                                  tensorflow/core/framework/tensor.proto:14-84
//   TensorShapeProto{ repeated Dim dim = 2 { size = 1; } }                         tensor_shape.proto:13-46
//   DataType        DT_FLOAT = 1, DT_INT64 = 9                                     types.proto:18,26
// DCNClient fills int64_val / float_val (DCNClient.java:101, :107); tensor_content (raw little-endian
// bytes, what most other clients send) is accepted too.
#pragma once
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

namespace ctr_wire {

struct Reader {
    const uint8_t* p; const uint8_t* end; bool ok = true;
    Reader(const uint8_t* b, size_t n) : p(b), end(b + n) {}
    bool done() const { return p >= end; }
    uint64_t varint() {
        uint64_t r = 0; int sh = 0;
        while (p < end && sh < 64) {
            const uint8_t b = *p++;
            r |= (uint64_t)(b & 0x7F) << sh;
            if (!(b & 0x80)) return r;
            sh += 7;
        }
        ok = false;
        return 0;
    }
    Reader sub() {                                   // length-delimited payload
        const uint64_t n = varint();
        if (!ok || n > (uint64_t)(end - p)) { ok = false; return Reader(p, 0); }
        Reader r(p, (size_t)n);
        p += n;
        return r;
    }
    void skip(uint32_t wire_type) {
        switch (wire_type) {
            case 0: varint(); break;
            case 1: if (end - p < 8) ok = false; else p += 8; break;
            case 2: sub(); break;
            case 5: if (end - p < 4) ok = false; else p += 4; break;
            default: ok = false;
        }
    }
};

struct Tensor {
    int dtype = 0;
    std::vector<int64_t> dims;
    std::vector<int64_t> i64;
    std::vector<float> f32;
};

inline bool parse_tensor(Reader r, Tensor* t) {
    std::string content;
    while (r.ok && !r.done()) {
        const uint64_t tag = r.varint();
        const uint32_t field = (uint32_t)(tag >> 3), wt = (uint32_t)(tag & 7);
        if (field == 1 && wt == 0) t->dtype = (int)r.varint();
        else if (field == 2 && wt == 2) {
            Reader s = r.sub();
            while (s.ok && !s.done()) {
                const uint64_t t2 = s.varint();
                if ((t2 >> 3) == 2 && (t2 & 7) == 2) {
                    Reader d = s.sub();
                    int64_t size = 0;
                    while (d.ok && !d.done()) {
                        const uint64_t t3 = d.varint();
                        if ((t3 >> 3) == 1 && (t3 & 7) == 0) size = (int64_t)d.varint(); else d.skip((uint32_t)(t3 & 7));
                    }
                    if (!d.ok) return false;
                    t->dims.push_back(size);
                } else s.skip((uint32_t)(t2 & 7));
            }
            if (!s.ok) return false;
        } else if (field == 4 && wt == 2) {
            Reader c = r.sub();
            content.assign(reinterpret_cast<const char*>(c.p), (size_t)(c.end - c.p));
        } else if (field == 5 && wt == 2) {                       // packed float_val
            Reader c = r.sub();
            const size_t n = (size_t)(c.end - c.p) / 4;
            const size_t old = t->f32.size();
            t->f32.resize(old + n);
            memcpy(t->f32.data() + old, c.p, n * 4);
        } else if (field == 5 && wt == 5) {                       // unpacked float_val
            if (r.end - r.p < 4) return false;
            float v; memcpy(&v, r.p, 4); r.p += 4; t->f32.push_back(v);
        } else if (field == 10 && wt == 2) {                      // packed int64_val
            Reader c = r.sub();
            while (c.ok && !c.done()) t->i64.push_back((int64_t)c.varint());
            if (!c.ok) return false;
        } else if (field == 10 && wt == 0) {
            t->i64.push_back((int64_t)r.varint());
        } else r.skip(wt);
    }
    if (!r.ok) return false;
    if (!content.empty()) {
        if (t->dtype == 9) { t->i64.resize(content.size() / 8); memcpy(t->i64.data(), content.data(), t->i64.size() * 8); }
        else if (t->dtype == 1) { t->f32.resize(content.size() / 4); memcpy(t->f32.data(), content.data(), t->f32.size() * 4); }
    }
    return true;
}

// Fills ids / wts from a PredictRequest; returns false with `why` on a malformed or inconsistent request
// (what TF-Serving answers with INVALID_ARGUMENT).
inline bool parse_predict_request(const uint8_t* buf, size_t len, int field_num, std::vector<int64_t>* ids, std::vector<float>* wts,
                                  std::string* model_name, std::string* signature, std::string* why) {
    Reader r(buf, len);
    Tensor t_ids, t_wts;
    bool have_ids = false, have_wts = false;
    while (r.ok && !r.done()) {
        const uint64_t tag = r.varint();
        const uint32_t field = (uint32_t)(tag >> 3), wt = (uint32_t)(tag & 7);
        if (field == 1 && wt == 2) {                              // model_spec
            Reader m = r.sub();
            while (m.ok && !m.done()) {
                const uint64_t t2 = m.varint();
                if ((t2 & 7) == 2 && ((t2 >> 3) == 1 || (t2 >> 3) == 3)) {
                    Reader sv = m.sub();
                    ((t2 >> 3) == 1 ? model_name : signature)->assign(reinterpret_cast<const char*>(sv.p), (size_t)(sv.end - sv.p));
                } else m.skip((uint32_t)(t2 & 7));
            }
            if (!m.ok) { *why = "malformed model_spec"; return false; }
        } else if (field == 2 && wt == 2) {                       // inputs map entry
            Reader e = r.sub();
            std::string key;
            Tensor t;
            bool have_val = false;
            while (e.ok && !e.done()) {
                const uint64_t t2 = e.varint();
                if ((t2 >> 3) == 1 && (t2 & 7) == 2) { Reader k = e.sub(); key.assign(reinterpret_cast<const char*>(k.p), (size_t)(k.end - k.p)); }
                else if ((t2 >> 3) == 2 && (t2 & 7) == 2) { if (!parse_tensor(e.sub(), &t)) { *why = "malformed TensorProto"; return false; } have_val = true; }
                else e.skip((uint32_t)(t2 & 7));
            }
            if (!e.ok) { *why = "malformed inputs entry"; return false; }
            if (have_val && key == "feat_ids") { t_ids = std::move(t); have_ids = true; }
            else if (have_val && key == "feat_wts") { t_wts = std::move(t); have_wts = true; }
        } else r.skip(wt);
    }
    if (!r.ok) { *why = "malformed PredictRequest"; return false; }
    if (!have_ids || !have_wts) { *why = "inputs feat_ids and feat_wts are both required"; return false; }
    if (t_ids.dtype != 9 || t_wts.dtype != 1) { *why = "feat_ids must be DT_INT64 and feat_wts DT_FLOAT"; return false; }
    auto count = [](const Tensor& t) { int64_t c = 1; for (int64_t d : t.dims) c *= d; return t.dims.empty() ? (int64_t)-1 : c; };
    if (t_ids.dims.size() != 2 || t_ids.dims != t_wts.dims || t_ids.dims[1] != field_num) { *why = "both inputs need shape [n, FIELD_NUM]"; return false; }
    if (count(t_ids) != (int64_t)t_ids.i64.size() || count(t_wts) != (int64_t)t_wts.f32.size()) { *why = "declared shape disagrees with the value count"; return false; }
    *ids = std::move(t_ids.i64);
    *wts = std::move(t_wts.f32);
    return true;
}

inline void put_varint(std::string* s, uint64_t v) {
    while (v >= 0x80) { s->push_back((char)(v | 0x80)); v >>= 7; }
    s->push_back((char)v);
}
inline void put_tag(std::string* s, uint32_t field, uint32_t wt) { put_varint(s, ((uint64_t)field << 3) | wt); }
inline void put_bytes(std::string* s, uint32_t field, const std::string& b) { put_tag(s, field, 2); put_varint(s, b.size()); s->append(b); }

// PredictResponse { outputs["prediction_node"] = TensorProto{DT_FLOAT, shape [n,1], float_val}, model_spec }
inline std::string build_predict_response(const float* scores, int64_t n, const std::string& model_name, const std::string& signature) {
    std::string shape;
    for (int64_t d : {n, (int64_t)1}) {
        std::string dim;
        put_tag(&dim, 1, 0); put_varint(&dim, (uint64_t)d);
        put_bytes(&shape, 2, dim);
    }
    std::string tensor;
    put_tag(&tensor, 1, 0); put_varint(&tensor, 1);                                   // DT_FLOAT
    put_bytes(&tensor, 2, shape);
    put_tag(&tensor, 5, 2); put_varint(&tensor, (uint64_t)n * 4);
    tensor.append(reinterpret_cast<const char*>(scores), (size_t)n * 4);
    std::string entry;
    put_bytes(&entry, 1, "prediction_node");
    put_bytes(&entry, 2, tensor);
    std::string out;
    put_bytes(&out, 1, entry);
    std::string spec;
    if (!model_name.empty()) put_bytes(&spec, 1, model_name);
    if (!signature.empty()) put_bytes(&spec, 3, signature);
    if (!spec.empty()) put_bytes(&out, 2, spec);
    return out;
}

}  // namespace ctr_wire

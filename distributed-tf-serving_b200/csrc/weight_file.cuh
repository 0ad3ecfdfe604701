// weight_file.cuh -- reader of the CTRW0001 weight file (format: DESIGN.md section 4, writer: weights.py).
#pragma once
namespace {
// ------------------------------------------------------------------ weight file (format: DESIGN.md)
constexpr char kMagic[8] = {'C', 'T', 'R', 'W', '0', '0', '0', '1'};
constexpr size_t kHeaderBytes = 128;
constexpr uint64_t kTagE = 1, kTagW1 = 2;
constexpr int kMaxHiddenWidth = 16384;

struct ModelDesc {
    int32_t kind = 0, F = 0, K = 0, id_mode = 0, table_dtype = 0, table_mode = 0, n_hidden = 0;
    int32_t hidden[CTR_MAX_HIDDEN] = {0, 0, 0, 0};
    int32_t n_cross = 0;
    int64_t V = 0;
    uint64_t seed = 0;
    float scale_e = 0, scale_w1 = 0;
    int d() const { return F * K; }
    int d_pad() const { return round_up(F * K, 128); }
    int head_in() const { return kind == CTR_MODEL_LR ? 0 : (n_hidden ? hidden[n_hidden - 1] : 0) + (kind == CTR_MODEL_DCN_V2 ? d() : 0); }
};

struct HostWeights {
    ModelDesc m;
    std::vector<float> blob;        // everything after the header
    size_t off_b0 = 0, off_w1 = 0, off_emb = 0, off_wout = 0, off_bout = 0;
    std::vector<size_t> off_mlp_w, off_mlp_b, off_cross_w, off_cross_b;
};

int32_t read_weight_file(const char* path, HostWeights* hw) {
    FILE* f = fopen(path, "rb");
    if (!f) return fail(CTR_ERR_IO, std::string("cannot open weight file ") + path);
    unsigned char hdr[kHeaderBytes];
    if (fread(hdr, 1, kHeaderBytes, f) != kHeaderBytes || memcmp(hdr, kMagic, 8) != 0) {
        fclose(f);
        return fail(CTR_ERR_IO, "weight file: bad header / magic");
    }
    ModelDesc& m = hw->m;
    int32_t iv[12];
    memcpy(iv, hdr + 8, sizeof(iv));
    m.kind = iv[0]; m.F = iv[1]; m.K = iv[2]; m.id_mode = iv[3]; m.table_dtype = iv[4]; m.table_mode = iv[5]; m.n_hidden = iv[6];
    for (int i = 0; i < CTR_MAX_HIDDEN; ++i) m.hidden[i] = iv[7 + i];
    m.n_cross = iv[11];
    memcpy(&m.V, hdr + 56, 8);
    memcpy(&m.seed, hdr + 64, 8);
    memcpy(&m.scale_e, hdr + 72, 4);
    memcpy(&m.scale_w1, hdr + 76, 4);

    auto bad = [&](const char* why) { fclose(f); return fail(CTR_ERR_IO, std::string("weight file: ") + why); };
    if (m.kind < 0 || m.kind > 3) return bad("unknown model kind");
    if (m.K != kEmbDim) return bad("embedding dim must be 16");
    if (m.F < 1 || m.F > CTR_MAX_FIELDS) return bad("field count out of range (1..48)");
    if (m.V < 1 || m.V > 0xFFFFFFFEll) return bad("row count out of range (1 .. 2^32 - 2: the kernels index rows with 32 bits)");
    if (m.id_mode != CTR_ID_DIRECT && m.id_mode != CTR_ID_HASHED) return bad("unknown id mode");
    if (m.table_dtype != CTR_TABLE_FP32 && m.table_dtype != CTR_TABLE_F16) return bad("unknown table dtype");
    if (m.table_mode != 0 && m.table_mode != 1) return bad("unknown table mode (0 = raw rows in the file, 1 = procedural)");
    if (m.id_mode == CTR_ID_HASHED && (m.V & (m.V - 1))) return bad("hashed tables need a power-of-two row count");
    if (m.n_hidden < 0 || m.n_hidden > CTR_MAX_HIDDEN) return bad("too many hidden layers");
    if (m.kind != CTR_MODEL_LR && m.n_hidden < 1) return bad("tower needs at least one hidden layer");
    for (int i = 0; i < m.n_hidden; ++i)
        if (m.hidden[i] < 8 || m.hidden[i] % 8 || m.hidden[i] > kMaxHiddenWidth) return bad("hidden sizes must be multiples of 8 in 8..16384");
    if (m.kind != CTR_MODEL_DCN_V2) m.n_cross = 0;
    if (m.n_cross < 0 || m.n_cross > 4) return bad("cross layer count out of range (0..4)");
    if (m.kind == CTR_MODEL_DCN_V2 && m.n_cross < 1) return bad("DCN-v2 needs at least one cross layer");

    size_t n = 1;
    hw->off_b0 = 0;
    if (m.table_mode == 0) { hw->off_w1 = n; n += (size_t)m.V; hw->off_emb = n; n += (size_t)m.V * m.K; }
    if (m.kind != CTR_MODEL_LR) {
        int fan_in = m.d();
        for (int l = 0; l < m.n_hidden; ++l) {
            hw->off_mlp_w.push_back(n); n += (size_t)m.hidden[l] * fan_in;
            hw->off_mlp_b.push_back(n); n += (size_t)m.hidden[l];
            fan_in = m.hidden[l];
        }
        for (int l = 0; l < m.n_cross; ++l) {
            hw->off_cross_w.push_back(n); n += (size_t)m.d() * m.d();
            hw->off_cross_b.push_back(n); n += (size_t)m.d();
        }
        hw->off_wout = n; n += (size_t)m.head_in();
        hw->off_bout = n; n += 1;
    }
    // the header decides how many floats follow: check it against the file before allocating for it
    const long data_pos = ftell(f);
    if (data_pos < 0 || fseek(f, 0, SEEK_END) != 0) return bad("cannot seek");
    const long end_pos = ftell(f);
    if (end_pos < data_pos || fseek(f, data_pos, SEEK_SET) != 0) return bad("cannot seek");
    if ((uint64_t)(end_pos - data_pos) != (uint64_t)n * sizeof(float))
        return bad(end_pos - data_pos < (long)(n * sizeof(float)) ? "truncated tensor data" : "trailing bytes after the last tensor");
    try {
        hw->blob.resize(n);
    } catch (const std::exception&) {
        return bad("not enough host memory for the tensors the header describes");
    }
    size_t got = fread(hw->blob.data(), sizeof(float), n, f);
    unsigned char extra;
    bool trailing = fread(&extra, 1, 1, f) == 1;
    fclose(f);
    if (got != n) return fail(CTR_ERR_IO, "weight file: truncated tensor data");
    if (trailing) return fail(CTR_ERR_IO, "weight file: trailing bytes after the last tensor");
    return CTR_OK;
}

}  // namespace

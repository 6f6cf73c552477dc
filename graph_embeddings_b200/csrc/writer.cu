// writer.cu — EmbeddingTextWriter.write served from the device store.
//
// Replaces J/util/write/EmbeddingTextWriter.java:73-149: for every vertex that passes the
// node-type filter (:102-109) and the URI-prefix filter (:114-124), one line of
// String.format("%11.6E", v) joined by tabs in <name>.vectors.tsv (:133-136) and one line
// `key \t TYPE` in <name>.dict.tsv (:137-145); both files start with the '#' settings block
// (:40-70), which the host formats and passes in as `header`.
//
// Rows are produced by extract_avg ((focus+context)/2 in f32, Optimizer.java:130-140) in chunks,
// copied to pinned host memory and formatted by a pool of host threads; nothing of size V*dim
// is ever materialised as double[] (10 GB on the Java heap at config 4).
#include "common.cuh"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <thread>

// ------------------------------------------------------------------------------------------
// java.util.Formatter "%11.6E" for a double.  Java does not round the exact binary value: it
// takes the SHORTEST decimal digit string that round-trips (FloatingDecimal, as Double.toString)
// and rounds THAT half-up to 6 decimals (FormattedFloatingDecimal.applyPrecision).  C's printf
// rounds the exact value half-even, which differs when the shortest digits end in a 5 exactly at
// the cut.  Fast path: 17 significant digits decide unless the tail is a near-tie; slow path:
// find the shortest round-trip string.  Output: [-]d.ddddddE[+-]dd (exponent >= 2 digits), left
// padded to width 11 (never needed: the shortest form has 12 characters).  NaN / Infinity print
// as Java does.
// ------------------------------------------------------------------------------------------
static int format_java_e(double v, char* out) {
    if (v != v) { strcpy(out, "        NaN"); return 11; }
    if (v == HUGE_VAL) { strcpy(out, "   Infinity"); return 11; }
    if (v == -HUGE_VAL) { strcpy(out, "  -Infinity"); return 11; }
    char buf[40];
    // "%.16E": [-]d.dddddddddddddddd E[+-]dd  (17 significant digits)
    snprintf(buf, sizeof buf, "%.16E", v);
    const char* p = buf;
    const bool neg = *p == '-';
    if (neg) p++;
    char dig[20];
    int nd = 0;
    dig[nd++] = p[0];
    for (int k = 2; k < 18; k++) dig[nd++] = p[k];  // 16 fraction digits
    int ex = atoi(p + 19);                           // after 'E'
    // near-tie test on digits 8..17 (index 7..16): "5000000000", "4999999999" or "5000000001"-like
    bool tie = false;
    if (dig[7] == '5') { tie = true; for (int k = 8; k < 16; k++) if (dig[k] != '0') { tie = false; break; } }
    else if (dig[7] == '4') { tie = true; for (int k = 8; k < 16; k++) if (dig[k] != '9') { tie = false; break; } }
    if (tie) {
        // shortest round-trip digits, as Double.toString would give them
        for (int prec = 0; prec <= 16; prec++) {
            char t[40];
            snprintf(t, sizeof t, "%.*E", prec, v);
            if (strtod(t, nullptr) == v) {
                const char* q = t + (neg ? 1 : 0);
                nd = 0;
                dig[nd++] = q[0];
                int k = 1;
                if (q[k] == '.') { k++; while (q[k] != 'E') dig[nd++] = q[k++]; }
                while (q[k] != 'E') k++;
                ex = atoi(q + k + 1);
                while (nd < 17) dig[nd++] = '0';
                break;
            }
        }
    }
    // half-up at 7 significant digits
    int d7[8];
    for (int k = 0; k < 7; k++) d7[k] = dig[k] - '0';
    if (dig[7] >= '5') {
        int k = 6;
        while (k >= 0 && ++d7[k] == 10) { d7[k] = 0; k--; }
        if (k < 0) { d7[0] = 1; for (int j = 1; j < 7; j++) d7[j] = 0; ex++; }
    }
    if (v == 0.0) ex = 0;
    char* o = out;
    if (neg) *o++ = '-';
    *o++ = (char)('0' + d7[0]);
    *o++ = '.';
    for (int k = 1; k < 7; k++) *o++ = (char)('0' + d7[k]);
    *o++ = 'E';
    *o++ = ex < 0 ? '-' : '+';
    const int ae = ex < 0 ? -ex : ex;
    if (ae >= 100) { *o++ = (char)('0' + ae / 100); }
    *o++ = (char)('0' + (ae / 10) % 10);
    *o++ = (char)('0' + ae % 10);
    *o = 0;
    return (int)(o - out);
}

extern "C" int ge_format_e(double v, char* out, int32_t cap) {
    char tmp[48];
    const int n = format_java_e(v, tmp);
    if (!out || cap <= n) return -n;
    memcpy(out, tmp, (size_t)n + 1);
    return n;
}

namespace {
template <typename T>
__global__ void extract_rows(const float* __restrict__ W, const float* __restrict__ C, int64_t row0, int32_t rows,
                             int32_t dim, int32_t S, T* __restrict__ out) {
    const int64_t total = (int64_t)rows * dim, stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < total; k += stride) {
        const int64_t r = k / dim;
        const int32_t d = (int32_t)(k - r * dim);
        out[k] = (T)__fdiv_rn(__fadd_rn(W[(row0 + r) * S + d], C[(row0 + r) * S + d]), 2.0f);
    }
}

const char* const TYPE_NAME[3] = {"URI", "BLANK", "LITERAL"};  // NodeInfo (J/convert/util/NodeInfo.java:8-34)
}  // namespace

extern "C" int ge_write_embeddings(ge_ctx* ctx, const char* vectors_path, const char* dict_path, const char* header,
                                   const char* const* keys, const uint8_t* types, const uint8_t* type_enabled,
                                   const char* const* prefixes, const int32_t* nprefix, int64_t* written) {
    if (!ctx || !vectors_path || !dict_path) return GE_ERR_INVALID;
    if (!ctx->have_store) { ctx->err = "no embedding store"; return GE_ERR_STATE; }
    if (ctx->dist) { ctx->err = "ge_write_embeddings: single-GPU store only (use ge_dist_extract_f32 per rank)"; return GE_ERR_STATE; }
    cudaSetDevice(ctx->cfg.device);
    const int32_t V = ctx->V, dim = ctx->cfg.dim;
    FILE* fv = fopen(vectors_path, "wb");
    FILE* fd = fopen(dict_path, "wb");
    if (!fv || !fd) {
        if (fv) fclose(fv);
        if (fd) fclose(fd);
        ctx->err = std::string("cannot open output file: ") + (fv ? dict_path : vectors_path);
        return GE_ERR_INVALID;
    }
    if (header) { fputs(header, fd); fputs(header, fv); }   // writeConfig(dict); writeConfig(vect)  :92-93
    fputs("key\ttype\n", fd);                                // :95

    // prefix lists per node type: prefixes[] is the concatenation URI | BLANK | LITERAL
    const char* const* pl[3] = {nullptr, nullptr, nullptr};
    int32_t np[3] = {0, 0, 0};
    if (prefixes && nprefix) {
        int32_t at = 0;
        for (int t = 0; t < 3; t++) { pl[t] = prefixes + at; np[t] = nprefix[t]; at += nprefix[t]; }
    }

    const int32_t chunk = 16384;
    float *d_out = nullptr, *h_out = nullptr;
    cudaError_t e = cudaMalloc(&d_out, (size_t)chunk * dim * sizeof(float));
    if (e == cudaSuccess) e = cudaMallocHost(&h_out, (size_t)chunk * dim * sizeof(float));
    if (e != cudaSuccess) { cudaFree(d_out); fclose(fv); fclose(fd); GE_CUDA(e); }
    unsigned nthr = std::thread::hardware_concurrency();
    if (nthr < 1) nthr = 1;
    if (nthr > 32) nthr = 32;
    std::vector<std::string> vbuf(nthr), dbuf(nthr);
    std::vector<int64_t> cnt(nthr);
    int64_t total = 0;
    int rc = GE_OK;
    for (int64_t row0 = 0; row0 < V && rc == GE_OK; row0 += chunk) {
        const int32_t rows = (int32_t)((V - row0) < chunk ? (V - row0) : chunk);
        extract_rows<float><<<ctx->sm_count * 4, 256, 0, ctx->stream>>>(ctx->d_W, ctx->d_C, row0, rows, dim, ctx->S, d_out);
        e = cudaMemcpyAsync(h_out, d_out, (size_t)rows * dim * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess) { ctx->err = std::string("writer: ") + cudaGetErrorString(e); rc = GE_ERR_CUDA; break; }
        auto work = [&](unsigned t) {
            std::string& vb = vbuf[t]; std::string& db = dbuf[t];
            vb.clear(); db.clear(); cnt[t] = 0;
            const int32_t lo = (int32_t)((int64_t)rows * t / nthr), hi = (int32_t)((int64_t)rows * (t + 1) / nthr);
            char num[48], idx[24];
            for (int32_t r = lo; r < hi; r++) {
                const int64_t i = row0 + r;
                const int type = types ? types[i] : 0;
                if (type > 2 || (type_enabled && !type_enabled[type])) continue;   // :102-109
                const char* key = keys ? keys[i] : nullptr;
                if (!key) { snprintf(idx, sizeof idx, "%lld", (long long)i); key = idx; }
                if (np[type] > 0) {                                                // :114-124 noneMatch(key::startsWith)
                    bool any = false;
                    for (int32_t k = 0; k < np[type] && !any; k++) any = strncmp(key, pl[type][k], strlen(pl[type][k])) == 0;
                    if (!any) continue;
                }
                const float* row = h_out + (size_t)r * dim;
                for (int32_t d = 0; d < dim; d++) {                                // :133-136
                    const int n = format_java_e((double)row[d], num);
                    if (d) vb.push_back('\t');
                    vb.append(num, (size_t)n);
                }
                vb.push_back('\n');
                for (const char* q = key; *q; q++) if (*q != '\n' && *q != '\r' && *q != '\t') db.push_back(*q);  // :139-141
                db.push_back('\t');
                db.append(TYPE_NAME[type]);
                db.push_back('\n');
                cnt[t]++;
            }
        };
        std::vector<std::thread> th;
        for (unsigned t = 1; t < nthr; t++) th.emplace_back(work, t);
        work(0);
        for (auto& x : th) x.join();
        for (unsigned t = 0; t < nthr; t++) {
            if (fwrite(vbuf[t].data(), 1, vbuf[t].size(), fv) != vbuf[t].size() ||
                fwrite(dbuf[t].data(), 1, dbuf[t].size(), fd) != dbuf[t].size()) {
                ctx->err = "writer: short write"; rc = GE_ERR_INVALID; break;
            }
            total += cnt[t];
        }
    }
    cudaFree(d_out);
    cudaFreeHost(h_out);
    if (fclose(fv) != 0 || fclose(fd) != 0) { if (rc == GE_OK) { ctx->err = "writer: close failed"; rc = GE_ERR_INVALID; } }
    if (written) *written = total;
    return rc;
}

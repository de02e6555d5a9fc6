// Device-side ingestion of the reference's text formats: once the step runs in ~0.1 ms the Python
// parsers (linecache + split, FNN_wnzh.py:224-253; ipinyou.py:42-65; get_fxy, baseline.py:190-198) are >99 % of the wall
// time, so the raw file bytes are uploaded as they are and turned into the engines' [N, F] id / weight blocks on the GPU.
//
//   1. dctr_text_index_lines   byte offset of every non-empty line, in file order: per-4-KB-tile count of line starts,
//                              one-CTA exclusive scan over the tiles, ordered fill (three short kernels, two reads of the text)
//   2. dctr_text_parse_sparse  `y [skipped tokens] id:val id:val ...`    (train.fm.txt, *.yzx.txt)
//      dctr_text_parse_criteo  `y \t n_1..n_K \t c_1..c_C`               (Criteo indexed TSV)
//      A CTA owns 32 consecutive lines = one contiguous byte range, copied to shared memory with 128-bit loads; a warp then
//      walks a line 32 bytes at a time: a ballot marks the bytes that start a token, popc over the lower lanes gives each
//      start its token index, and the lane that holds a start converts that one number serially (<= 19 digits).
// Integer work: ids / offsets / counts are bit-exact with the reference parsers on well-formed lines.
#include "common.cuh"

namespace dctr {

constexpr int kIdxThreads = 256;
constexpr int kIdxBytesPerThread = 16;
constexpr int kIdxTile = kIdxThreads * kIdxBytesPerThread;     // 4096 bytes per CTA

// bit k of the result: byte p + k starts a line (not '\n' itself, preceded by '\n' or the start of the buffer)
__device__ __forceinline__ unsigned line_start_mask(const uint8_t* __restrict__ text, int64_t nbytes, int64_t p) {
    if (p >= nbytes) return 0u;
    uint8_t b[kIdxBytesPerThread];
    if (p + kIdxBytesPerThread <= nbytes) {
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(text + p));
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int k = 0; k < kIdxBytesPerThread; ++k) b[k] = (uint8_t)(w[k >> 2] >> (8 * (k & 3)));
    } else {
#pragma unroll
        for (int k = 0; k < kIdxBytesPerThread; ++k) b[k] = p + k < nbytes ? __ldg(text + p + k) : (uint8_t)'\n';
    }
    uint8_t prev = p > 0 ? __ldg(text + p - 1) : (uint8_t)'\n';
    unsigned m = 0;
#pragma unroll
    for (int k = 0; k < kIdxBytesPerThread; ++k) {
        m |= (unsigned)(b[k] != '\n' && prev == '\n' && p + k < nbytes) << k;
        prev = b[k];
    }
    return m;
}

// exclusive scan of one int per thread over the CTA (kIdxThreads threads); *total = CTA sum
__device__ __forceinline__ int cta_exclusive_scan(int v, int* total) {
    __shared__ int warp_sums[kIdxThreads / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) warp_sums[warp] = inc;
    __syncthreads();
    int before = 0, all = 0;
#pragma unroll
    for (int w = 0; w < kIdxThreads / 32; ++w) {
        const int s = warp_sums[w];
        if (w < warp) before += s;
        all += s;
    }
    *total = all;
    return before + inc - v;
}

__global__ void __launch_bounds__(kIdxThreads)
text_count_kernel(const uint8_t* __restrict__ text, int64_t nbytes, int32_t* __restrict__ tile_count) {
    pdl_sync();
    const int64_t p = (int64_t)blockIdx.x * kIdxTile + (int64_t)threadIdx.x * kIdxBytesPerThread;
    int total;
    cta_exclusive_scan(__popc(line_start_mask(text, nbytes, p)), &total);
    if (threadIdx.x == 0) tile_count[blockIdx.x] = total;
}

// one CTA: tile_off[t] = sum of tile_count[0..t), n_lines = the grand total; writes the end sentinel when it fits
__global__ void __launch_bounds__(1024)
text_scan_kernel(const int32_t* __restrict__ tile_count, int64_t tiles, int64_t* __restrict__ tile_off,
                 int64_t* __restrict__ n_lines, int64_t* __restrict__ line_start, int64_t max_lines, int64_t nbytes) {
    pdl_sync();
    __shared__ int64_t part[1024];
    const int t = threadIdx.x;
    const int64_t per = (tiles + 1023) / 1024;
    const int64_t lo = (int64_t)t * per, hi = lo + per < tiles ? lo + per : tiles;
    int64_t s = 0;
    for (int64_t i = lo; i < hi; ++i) s += tile_count[i];
    part[t] = s;
    __syncthreads();
    if (t == 0) {                                  // 1024 serial adds: negligible next to the two passes over the text
        int64_t run = 0;
        for (int i = 0; i < 1024; ++i) { const int64_t v = part[i]; part[i] = run; run += v; }
        n_lines[0] = run;
        if (run <= max_lines) line_start[run] = nbytes;
    }
    __syncthreads();
    int64_t run = part[t];
    for (int64_t i = lo; i < hi; ++i) { tile_off[i] = run; run += tile_count[i]; }
}

__global__ void __launch_bounds__(kIdxThreads)
text_fill_kernel(const uint8_t* __restrict__ text, int64_t nbytes, const int64_t* __restrict__ tile_off,
                 int64_t* __restrict__ line_start, int64_t max_lines) {
    pdl_sync();
    const int64_t p = (int64_t)blockIdx.x * kIdxTile + (int64_t)threadIdx.x * kIdxBytesPerThread;
    unsigned m = line_start_mask(text, nbytes, p);
    int total;
    int64_t idx = tile_off[blockIdx.x] + cta_exclusive_scan(__popc(m), &total);
    while (m) {
        const int k = __ffs(m) - 1;
        m &= m - 1;
        if (idx <= max_lines) line_start[idx] = p + k;      // slot max_lines = end bound of the last line that is kept
        ++idx;
    }
}

// ---- parsing --------------------------------------------------------------------------------------------------------
constexpr int kParseWarps = 8;
constexpr int kParseThreads = kParseWarps * 32;
constexpr int kLinesPerWarp = 4;
constexpr int kLinesPerCta = kParseWarps * kLinesPerWarp;       // 32 consecutive lines = one contiguous byte range
constexpr int kStageBytes = 24 * 1024;

struct ParseArgs {
    const uint8_t* text; int64_t nbytes; const int64_t* line_start; const int64_t* n_lines; int64_t max_lines;
    // sparse: skip tokens after y, J feature slots per line, pad id
    int skip, J; int32_t pad_id;
    // criteo: n_num numeric + n_cat categorical fields, F = n_num + n_cat
    int n_num, n_cat; const int64_t* offsets; const double* max_vals;
    int32_t* ids; float* wts; float* y; int32_t* status;
};

__device__ __forceinline__ bool is_ws(uint8_t c) { return c == ' ' || (c >= 9 && c <= 13); }      // str.split() whitespace

struct LineBytes {            // the CTA's byte range, staged in shared memory when it fits
    const uint8_t* text; const uint8_t* sbuf; int64_t a0; bool staged;
    __device__ __forceinline__ uint8_t get(int64_t q) const { return staged ? sbuf[q - a0] : __ldg(text + q); }
};

// decimal number starting at byte q (optional '-', digits, with FRAC an optional '.digits'); returns the terminating byte
// in `term`, the digit count in `nd`
template <bool FRAC>
__device__ __forceinline__ double parse_number(const LineBytes& lb, int64_t q, int64_t le, int64_t* ival, int* nd, uint8_t* term) {
    uint8_t ch = lb.get(q);
    bool neg = false;
    if (ch == '-' || ch == '+') { neg = ch == '-'; ++q; ch = q < le ? lb.get(q) : (uint8_t)'\n'; }
    int64_t v = 0;
    int n = 0;
    while (ch >= '0' && ch <= '9') {
        if (n < 18) v = v * 10 + (ch - '0');
        ++n; ++q;
        ch = q < le ? lb.get(q) : (uint8_t)'\n';
    }
    double val = (double)v;
    if (FRAC && ch == '.') {
        ++q;
        ch = q < le ? lb.get(q) : (uint8_t)'\n';
        int64_t f = 0;
        double scale = 1.0;
        int nf = 0;
        while (ch >= '0' && ch <= '9') {
            if (nf < 18) { f = f * 10 + (ch - '0'); scale *= 10.0; }
            ++nf; ++n; ++q;
            ch = q < le ? lb.get(q) : (uint8_t)'\n';
        }
        val += (double)f / scale;
    }
    *ival = neg ? -v : v;
    *nd = n > 18 ? -1 : n;          // more than 18 digits: reported as malformed
    *term = ch;
    return neg ? -val : val;
}

template <bool CRITEO>
__global__ void __launch_bounds__(kParseThreads)
text_parse_kernel(const ParseArgs a) {
    pdl_sync();
    __shared__ __align__(16) uint8_t sbuf[kStageBytes];
    int64_t n = a.n_lines[0];
    if (n > a.max_lines) n = a.max_lines;
    const int64_t L0 = (int64_t)blockIdx.x * kLinesPerCta;
    if (L0 >= n) return;                                     // whole CTA: no barrier has been reached yet
    const int64_t L1 = L0 + kLinesPerCta < n ? L0 + kLinesPerCta : n;
    const int64_t b0 = a.line_start[L0], b1 = a.line_start[L1];
    LineBytes lb;
    lb.text = a.text; lb.sbuf = sbuf; lb.a0 = b0 & ~(int64_t)15;
    lb.staged = b1 - lb.a0 <= kStageBytes;
    if (lb.staged) {
        for (int64_t off = lb.a0 + (int64_t)threadIdx.x * 16; off < b1; off += (int64_t)kParseThreads * 16) {
            if (off + 16 <= a.nbytes) {
                *reinterpret_cast<uint4*>(sbuf + (off - lb.a0)) = __ldg(reinterpret_cast<const uint4*>(a.text + off));
            } else {
                for (int k = 0; k < 16 && off + k < a.nbytes; ++k) sbuf[off - lb.a0 + k] = __ldg(a.text + off + k);
            }
        }
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned lt = (1u << lane) - 1u;
    const int F = CRITEO ? a.n_num + a.n_cat : a.J;
    for (int li = 0; li < kLinesPerWarp; ++li) {
        const int64_t line = L0 + (int64_t)li * kParseWarps + warp;
        if (line >= L1) break;
        const int64_t ls = a.line_start[line], le = a.line_start[line + 1];
        int32_t* ids = a.ids + line * F;
        float* wts = a.wts ? a.wts + line * F : nullptr;
        int count = 0;                       // sparse: whitespace tokens so far; criteo: tabs so far
        bool bad = false;
        uint8_t prev = CRITEO ? (uint8_t)'\t' : (uint8_t)' ';
        bool done = false;                   // criteo: the line's '\n' has been passed
        for (int64_t base = ls; base < le && !done; base += 32) {
            const int64_t q = base + lane;
            const uint8_t c = q < le ? lb.get(q) : (uint8_t)'\n';
            uint8_t pc = (uint8_t)__shfl_up_sync(0xffffffffu, (unsigned)c, 1);
            if (lane == 0) pc = prev;
            if (CRITEO) {
                const unsigned nl = __ballot_sync(0xffffffffu, c == '\n');
                const unsigned before_nl = nl ? ((1u << (__ffs(nl) - 1)) - 1u) | (1u << (__ffs(nl) - 1)) : 0xffffffffu;   // up to and incl. the '\n'
                const bool live = (before_nl >> lane) & 1u;
                const unsigned tabs = __ballot_sync(0xffffffffu, live && c == '\t');
                const bool start = live && q < le && pc == '\t';
                const int fi = count + __popc(tabs & lt);
                if (start) {
                    int64_t iv; int nd; uint8_t term;
                    const double val = parse_number<true>(lb, q, le, &iv, &nd, &term);
                    if (nd <= 0 || !(term == '\t' || term == '\n' || term == '\r')) bad = true;
                    if (fi == 0) {
                        a.y[line] = (float)val;
                    } else if (fi <= a.n_num) {                  // _x = float(field) / max_vals[i-1], id = i-1 (baseline.py:194-196)
                        ids[fi - 1] = fi - 1;
                        if (wts) wts[fi - 1] = (float)(val / a.max_vals[fi - 1]);
                    } else if (fi <= F) {                        // int(cat) + offsets[i] (baseline.py:195), weight 1
                        const int j = fi - 1 - a.n_num;
                        const int64_t g = iv + a.offsets[j];
                        if (g < 0 || g > 0x7fffffffLL) bad = true;
                        ids[fi - 1] = (int32_t)g;
                        if (wts) wts[fi - 1] = 1.f;
                    }
                }
                count += __popc(tabs);
                done = nl != 0;
            } else {
                const bool start = q < le && !is_ws(c) && is_ws(pc);
                const unsigned sm = __ballot_sync(0xffffffffu, start);
                const int w = count + __popc(sm & lt);
                if (start) {
                    int64_t iv; int nd; uint8_t term;
                    parse_number<false>(lb, q, le, &iv, &nd, &term);
                    if (nd <= 0 || !(is_ws(term) || (term == ':' && w > 0))) bad = true;
                    if (w == 0) {
                        a.y[line] = (float)iv;                   // y = int(s[0]) (FNN_wnzh.py:250, ipinyou.py:52)
                    } else if (w > a.skip) {                     // the token before the first ':' is the feature id
                        const int j = w - 1 - a.skip;
                        if (iv < 0 || iv > 0x7fffffffLL) bad = true;
                        if (j < a.J) {
                            ids[j] = (int32_t)iv;
                            if (wts) wts[j] = 1.f;               // values are forced to 1 (ipinyou.py:55) / ignored (FNN_wnzh.py:251)
                        }
                    }
                }
                count += __popc(sm);
            }
            prev = (uint8_t)__shfl_sync(0xffffffffu, (unsigned)c, 31);
        }
        bad = __any_sync(0xffffffffu, bad);
        // slots the line did not fill
        int filled, stat;
        if (CRITEO) {
            filled = count;                                    // fields = tabs + 1, the first one is y
            stat = count + 1;
        } else {
            filled = count - 1 - a.skip;
            if (filled < 0) filled = 0;
            stat = count == 0 ? -1 : filled;                   // whitespace-only line: -1
            if (count == 0 && lane == 0) a.y[line] = 0.f;
        }
        for (int j = filled + lane; j < F; j += 32) {
            ids[j] = CRITEO ? 0 : a.pad_id;
            if (wts) wts[j] = 0.f;
        }
        if (lane == 0) a.status[line] = bad ? -2 : stat;
    }
}

}  // namespace dctr

using namespace dctr;

static int64_t idx_tiles(int64_t nbytes) { return (nbytes + kIdxTile - 1) / kIdxTile; }

extern "C" size_t dctr_text_index_workspace(int64_t nbytes) {
    if (nbytes < 0) nbytes = 0;
    const int64_t tiles = idx_tiles(nbytes) + 1;
    return (size_t)tiles * (sizeof(int32_t) + sizeof(int64_t)) + 64;
}

extern "C" int dctr_text_index_lines(const uint8_t* text, int64_t nbytes, int64_t* line_start, int64_t max_lines,
                                     int64_t* n_lines, void* workspace, size_t workspace_bytes, void* stream) {
    DCTR_REQUIRE(nbytes >= 0 && max_lines >= 0 && line_start && n_lines, "text_index_lines: bad arguments");
    DCTR_REQUIRE(nbytes == 0 || (text && workspace), "text_index_lines: null pointer");
    DCTR_REQUIRE(((uintptr_t)text & 15) == 0 && ((uintptr_t)workspace & 7) == 0, "text_index_lines: text must be 16-byte, workspace 8-byte aligned");
    DCTR_REQUIRE(idx_tiles(nbytes) < ((int64_t)1 << 31), "text_index_lines: buffer too large");
    if (workspace_bytes < dctr_text_index_workspace(nbytes)) {
        set_error("text_index_lines: workspace %zu < %zu bytes", workspace_bytes, dctr_text_index_workspace(nbytes));
        return DCTR_EWORKSPACE;
    }
    cudaStream_t st = as_stream(stream);
    const int64_t tiles = idx_tiles(nbytes);
    int64_t* tile_off = static_cast<int64_t*>(workspace);
    int32_t* tile_count = reinterpret_cast<int32_t*>(tile_off + tiles + 1);
    int rc;
    if (tiles > 0) {
        launch_k(text_count_kernel, (int)tiles, kIdxThreads, 0, st, text, nbytes, tile_count);
        if ((rc = check_launch("text_count_kernel"))) return rc;
    }
    launch_k(text_scan_kernel, 1, 1024, 0, st, (const int32_t*)tile_count, tiles, tile_off, n_lines, line_start, max_lines, nbytes);
    if ((rc = check_launch("text_scan_kernel"))) return rc;
    if (tiles > 0) {
        launch_k(text_fill_kernel, (int)tiles, kIdxThreads, 0, st, text, nbytes, (const int64_t*)tile_off, line_start, max_lines);
        if ((rc = check_launch("text_fill_kernel"))) return rc;
    }
    return DCTR_OK;
}

extern "C" int dctr_text_parse_sparse(const uint8_t* text, int64_t nbytes, const int64_t* line_start, const int64_t* n_lines,
                                      int64_t max_lines, int skip, int J, int32_t pad_id, int32_t* feats, float* wts, float* y,
                                      int32_t* status, void* stream) {
    DCTR_REQUIRE(nbytes >= 0 && max_lines >= 0 && skip >= 0 && J > 0, "text_parse_sparse: bad shape");
    if (max_lines == 0 || nbytes == 0) return DCTR_OK;
    DCTR_REQUIRE(text && line_start && n_lines && feats && y && status, "text_parse_sparse: null pointer");
    DCTR_REQUIRE(((uintptr_t)text & 15) == 0, "text_parse_sparse: text must be 16-byte aligned");
    ParseArgs a{};
    a.text = text; a.nbytes = nbytes; a.line_start = line_start; a.n_lines = n_lines; a.max_lines = max_lines;
    a.skip = skip; a.J = J; a.pad_id = pad_id; a.ids = feats; a.wts = wts; a.y = y; a.status = status;
    launch_k(text_parse_kernel<false>, ceil_div(max_lines, kLinesPerCta), kParseThreads, 0, as_stream(stream), a);
    return check_launch("text_parse_kernel<sparse>");
}

extern "C" int dctr_text_parse_criteo(const uint8_t* text, int64_t nbytes, const int64_t* line_start, const int64_t* n_lines,
                                      int64_t max_lines, int n_num, int n_cat, const int64_t* offsets, const double* max_vals,
                                      int32_t* ids, float* wts, float* y, int32_t* status, void* stream) {
    DCTR_REQUIRE(nbytes >= 0 && max_lines >= 0 && n_num >= 0 && n_cat >= 0 && n_num + n_cat > 0, "text_parse_criteo: bad shape");
    if (max_lines == 0 || nbytes == 0) return DCTR_OK;
    DCTR_REQUIRE(text && line_start && n_lines && ids && y && status && (n_cat == 0 || offsets) && (n_num == 0 || max_vals),
                 "text_parse_criteo: null pointer");
    DCTR_REQUIRE(((uintptr_t)text & 15) == 0, "text_parse_criteo: text must be 16-byte aligned");
    ParseArgs a{};
    a.text = text; a.nbytes = nbytes; a.line_start = line_start; a.n_lines = n_lines; a.max_lines = max_lines;
    a.n_num = n_num; a.n_cat = n_cat; a.offsets = offsets; a.max_vals = max_vals; a.ids = ids; a.wts = wts; a.y = y; a.status = status;
    launch_k(text_parse_kernel<true>, ceil_div(max_lines, kLinesPerCta), kParseThreads, 0, as_stream(stream), a);
    return check_launch("text_parse_kernel<criteo>");
}

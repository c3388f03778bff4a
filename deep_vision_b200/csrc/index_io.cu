// index_io.cu — persistent binary sidecar of a resident chunk index.
// The reference writes `chunks.json` next to `full_content.md` (web/server.py:31986-31999) and re-parses + re-lowers
// every chunk on every prompt build (web/server.py:32028, 32075).  The sidecar ("chunks.dvbidx", same directory) holds
// the symbolised chunks and the sorted postings exactly as they sit in HBM, so a later process loads the index with
// plain reads + H2D copies instead of rebuilding it.  Keyed by a 32-byte caller tag (the host mirror uses a SHA-256 over
// content_sha256 / chunk_size / chunk_overlap / chunk count of the manifest): a stale sidecar is refused, never used.
//
// File (little-endian, every array starts on a 16-byte boundary):
//   header   "DVBIDX\1\0", u32 version, u32 n_segments, i64 n_chunks, n_bytes, n_post, n_keys, u8 tag[32], u32 sub_bits, pad to 128
//   segment  i64 chunk_base, n_chunks, n_bytes, n_post, n_keys, u64 checksum[7]  (128 B), then the 7 arrays
//            sym u16[n_bytes] | sym_off i64[n_chunks+1] | sym_cnt i32[n_chunks] | pc u32[n_post] | pv u32[n_post] |
//            ukeys u32[n_keys] | uoff i64[n_keys+1]
// Checksums are computed ON THE DEVICE (k_ix_checksum: position-weighted 64-bit sum) when saving and again after the
// upload when loading, so truncation, bit rot and a failed copy all surface as DVB_E_INVALID.
#include "index.cuh"
#include <stdlib.h>

#define IXIO_VERSION 1u
#define IXIO_SUB_BITS 2u                   // must equal IX_SUB_BITS of index.cu: the sort key layout is part of the format
#define IXIO_BUF (32u << 20)               // pinned bounce buffer (two of them: file I/O overlaps the PCIe copy)

static const char IXIO_MAGIC[8] = { 'D', 'V', 'B', 'I', 'D', 'X', 1, 0 };

struct IxioHeader {
    char magic[8]; uint32_t version, n_segs; int64_t n_chunks, n_bytes, n_post, n_keys; uint8_t tag[32]; uint32_t sub_bits; uint8_t pad[44];
};
struct IxioSegHeader { int64_t chunk_base, n_chunks, n_bytes, n_post, n_keys; uint64_t checksum[7]; uint8_t pad[32]; };
static_assert(sizeof(IxioHeader) == 128 && sizeof(IxioSegHeader) == 128, "sidecar headers are 128 bytes");

// sum over 16-bit units u_i of (u_i + 1) * (2 i + 1)  mod 2^64   (every array has an even byte count)
__global__ void __launch_bounds__(256) k_ix_checksum(const uint16_t* __restrict__ a, int64_t n16, unsigned long long* __restrict__ out) {
    unsigned long long acc = 0;
    const int64_t n64 = n16 >> 2;
    const bool aligned = (reinterpret_cast<uintptr_t>(a) & 7u) == 0;
    if (aligned) {
        const unsigned long long* w = reinterpret_cast<const unsigned long long*>(a);
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n64; i += (int64_t)gridDim.x * blockDim.x) {
            const unsigned long long v = w[i];
#pragma unroll
            for (int j = 0; j < 4; j++) acc += (((v >> (16 * j)) & 0xFFFFull) + 1ull) * (unsigned long long)(2 * (4 * i + j) + 1);
        }
    }
    for (int64_t i = (aligned ? n64 * 4 : 0) + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (int64_t)gridDim.x * blockDim.x)
        acc += ((unsigned long long)a[i] + 1ull) * (unsigned long long)(2 * i + 1);
    for (int d = 16; d > 0; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
    if ((threadIdx.x & 31) == 0 && acc) atomicAdd(out, acc);
}

// Structural validation of an uploaded segment: every index the query kernels will follow stays inside its array
// (a sidecar is data from disk; the checksums only prove it is the data that was written).
__global__ void __launch_bounds__(256) k_ix_validate(IxSeg sg, int* __restrict__ bad) {
    const int64_t t0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, step = (int64_t)gridDim.x * blockDim.x;
    bool b = false;
    for (int64_t c = t0; c < sg.n_chunks; c += step) {
        const int64_t o0 = sg.sym_off[c], o1 = sg.sym_off[c + 1];
        b |= o0 < 0 || o1 < o0 || o1 > sg.n_bytes || sg.sym_cnt[c] < 0 || sg.sym_cnt[c] > o1 - o0 || sg.sym_cnt[c] > 65535;
    }
    if (t0 == 0) b |= sg.sym_off[0] != 0 || sg.sym_off[sg.n_chunks] != sg.n_bytes;
    if (sg.n_post > 0) {
        for (int64_t k = t0; k < sg.n_keys; k += step) {
            const int64_t o0 = sg.uoff[k], o1 = sg.uoff[k + 1];
            b |= o0 < 0 || o1 <= o0 || o1 > sg.n_post || (k > 0 && sg.ukeys[k] <= sg.ukeys[k - 1]);
        }
        if (t0 == 0) b |= sg.uoff[0] != 0 || sg.uoff[sg.n_keys] != sg.n_post || sg.n_keys < 1;
        for (int64_t p = t0; p < sg.n_post; p += step) {
            const int64_t c = (int64_t)sg.pc[p];
            if (c >= sg.n_chunks) { b = true; continue; }
            b |= (int64_t)(sg.pv[p] & 0xFFFFu) + 1 >= (int64_t)sg.sym_cnt[c];                    // a pair needs position + 1 inside the chunk
        }
    }
    if (b) atomicExch(bad, 1);
}

namespace {
struct SegArrays { void* ptr[7]; size_t bytes[7]; };
SegArrays seg_arrays(const IxSeg& sg) {
    SegArrays a;
    void* p[7] = { sg.sym, sg.sym_off, sg.sym_cnt, sg.pc, sg.pv, sg.ukeys, sg.uoff };
    const size_t b[7] = { 2 * (size_t)sg.n_bytes, 8 * (size_t)(sg.n_chunks + 1), 4 * (size_t)sg.n_chunks, 4 * (size_t)sg.n_post, 4 * (size_t)sg.n_post,
                          sg.n_post > 0 ? 4 * (size_t)sg.n_keys : 0, sg.n_post > 0 ? 8 * (size_t)(sg.n_keys + 1) : 0 };
    for (int i = 0; i < 7; i++) { a.ptr[i] = p[i]; a.bytes[i] = sg.n_chunks > 0 ? b[i] : 0; }     // the empty index owns no arrays
    return a;
}

struct Bounce {
    uint8_t* buf[2] = { nullptr, nullptr };
    cudaEvent_t ev[2] = { nullptr, nullptr };
    unsigned long long* d_sum = nullptr;
    bool ok() const { return buf[0] && buf[1] && ev[0] && ev[1] && d_sum; }
    Bounce() {
        for (int i = 0; i < 2; i++) {
            if (cudaMallocHost((void**)&buf[i], IXIO_BUF) != cudaSuccess) buf[i] = nullptr;
            if (cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming) != cudaSuccess) ev[i] = nullptr;
        }
        if (cudaMalloc((void**)&d_sum, 8 * 8) != cudaSuccess) d_sum = nullptr;
        cudaGetLastError();
    }
    ~Bounce() {
        for (int i = 0; i < 2; i++) { if (buf[i]) cudaFreeHost(buf[i]); if (ev[i]) cudaEventDestroy(ev[i]); }
        if (d_sum) cudaFree(d_sum);
    }
};

bool pad16(FILE* f, size_t bytes) {
    static const uint8_t z[16] = { 0 };
    const size_t r = (16 - (bytes & 15)) & 15;
    return r == 0 || fwrite(z, 1, r, f) == r;
}
bool skip16(FILE* f, size_t bytes) {
    const size_t r = (16 - (bytes & 15)) & 15;
    return r == 0 || fseek(f, (long)r, SEEK_CUR) == 0;
}

// device -> file: the D2H copy of piece i+1 is in flight while piece i is written
bool write_array(FILE* f, const void* dptr, size_t bytes, Bounce& bb, cudaStream_t st) {
    const uint8_t* d = static_cast<const uint8_t*>(dptr);
    const size_t n_piece = (bytes + IXIO_BUF - 1) / IXIO_BUF;
    auto piece = [&](size_t i) { return std::min<size_t>(IXIO_BUF, bytes - i * IXIO_BUF); };
    if (n_piece) {
        if (cudaMemcpyAsync(bb.buf[0], d, piece(0), cudaMemcpyDeviceToHost, st) != cudaSuccess) return false;
        cudaEventRecord(bb.ev[0], st);
    }
    for (size_t i = 0; i < n_piece; i++) {
        if (i + 1 < n_piece) {
            if (cudaMemcpyAsync(bb.buf[(i + 1) & 1], d + (i + 1) * IXIO_BUF, piece(i + 1), cudaMemcpyDeviceToHost, st) != cudaSuccess) return false;
            cudaEventRecord(bb.ev[(i + 1) & 1], st);
        }
        if (cudaEventSynchronize(bb.ev[i & 1]) != cudaSuccess) return false;
        if (fwrite(bb.buf[i & 1], 1, piece(i), f) != piece(i)) return false;
    }
    return pad16(f, bytes);
}
// file -> device: piece i+1 is read from the file while piece i crosses PCIe
bool read_array(FILE* f, void* dptr, size_t bytes, Bounce& bb, cudaStream_t st) {
    uint8_t* d = static_cast<uint8_t*>(dptr);
    const size_t n_piece = (bytes + IXIO_BUF - 1) / IXIO_BUF;
    for (size_t i = 0; i < n_piece; i++) {
        const size_t m = std::min<size_t>(IXIO_BUF, bytes - i * IXIO_BUF);
        if (i >= 2 && cudaEventSynchronize(bb.ev[i & 1]) != cudaSuccess) return false;     // the copy that used this buffer is done
        if (fread(bb.buf[i & 1], 1, m, f) != m) return false;
        if (cudaMemcpyAsync(d + i * IXIO_BUF, bb.buf[i & 1], m, cudaMemcpyHostToDevice, st) != cudaSuccess) return false;
        cudaEventRecord(bb.ev[i & 1], st);
    }
    if (cudaStreamSynchronize(st) != cudaSuccess) return false;
    return skip16(f, bytes);
}

// checksums of the 7 arrays of a segment (one launch each, one D2H)
bool seg_checksums(dvb_ctx* ctx, const IxSeg& sg, Bounce& bb, cudaStream_t st, uint64_t out[7]) {
    const SegArrays a = seg_arrays(sg);
    if (cudaMemsetAsync(bb.d_sum, 0, 64, st) != cudaSuccess) return false;
    for (int i = 0; i < 7; i++) if (a.bytes[i]) {
        const int64_t n16 = (int64_t)(a.bytes[i] / 2);
        const int grid = (int)std::min<int64_t>((n16 / 4 + 255) / 256 + 1, (int64_t)ctx->num_sm * 8);
        k_ix_checksum<<<grid, 256, 0, st>>>(static_cast<const uint16_t*>(a.ptr[i]), n16, bb.d_sum + i);
        ctx->launches += 1;
    }
    unsigned long long h[8];
    if (cudaMemcpyAsync(h, bb.d_sum, 64, cudaMemcpyDeviceToHost, st) != cudaSuccess) return false;
    if (cudaStreamSynchronize(st) != cudaSuccess || cudaGetLastError() != cudaSuccess) return false;
    for (int i = 0; i < 7; i++) out[i] = h[i];
    return true;
}
}  // namespace

extern "C" int32_t dvb_chunk_index_save(dvb_ctx* ctx, const dvb_index* ix, const char* path, const uint8_t* tag32) {
    if (!ctx || !ix || !path || !*path || ix->device != ctx->device) { dvb_set_err(ctx, "dvb_chunk_index_save: invalid argument%s%s"); return DVB_E_INVALID; }
    dvb_scope sc(ctx);
    cudaStream_t st = sc.stream();
    Bounce bb;
    if (!bb.ok()) { dvb_set_err(ctx, "dvb_chunk_index_save: cannot allocate the pinned bounce buffers%s%s"); return DVB_E_NOMEM; }
    const std::string tmp = std::string(path) + ".tmp";              // written beside the target, renamed when complete
    FILE* f = fopen(tmp.c_str(), "wb");
    if (!f) { dvb_set_err(ctx, "dvb_chunk_index_save: cannot create %s", tmp.c_str()); return DVB_E_INVALID; }
    IxioHeader h; memset(&h, 0, sizeof(h));
    memcpy(h.magic, IXIO_MAGIC, 8); h.version = IXIO_VERSION; h.n_segs = (uint32_t)ix->segs.size();
    h.n_chunks = ix->n_chunks; h.n_bytes = ix->n_bytes; h.n_post = ix->n_post; h.n_keys = ix->n_keys; h.sub_bits = IXIO_SUB_BITS;
    if (tag32) memcpy(h.tag, tag32, 32);
    bool ok = fwrite(&h, 1, sizeof(h), f) == sizeof(h);
    for (size_t s = 0; ok && s < ix->segs.size(); s++) {
        const IxSeg& sg = ix->segs[s];
        IxioSegHeader sh; memset(&sh, 0, sizeof(sh));
        sh.chunk_base = sg.chunk_base; sh.n_chunks = sg.n_chunks; sh.n_bytes = sg.n_bytes; sh.n_post = sg.n_post; sh.n_keys = sg.n_keys;
        ok = seg_checksums(ctx, sg, bb, st, sh.checksum) && fwrite(&sh, 1, sizeof(sh), f) == sizeof(sh);
        const SegArrays a = seg_arrays(sg);
        for (int i = 0; ok && i < 7; i++) ok = write_array(f, a.ptr[i], a.bytes[i], bb, st);
    }
    ok = (fflush(f) == 0) && ok;
    ok = (fclose(f) == 0) && ok;
    cudaStreamSynchronize(st);
    if (ok && rename(tmp.c_str(), path) != 0) ok = false;
    if (!ok) { remove(tmp.c_str()); cudaGetLastError(); dvb_set_err(ctx, "dvb_chunk_index_save: write to %s failed", path); return DVB_E_INVALID; }
    return DVB_OK;
}

extern "C" int32_t dvb_chunk_index_load(dvb_ctx* ctx, const char* path, const uint8_t* tag32, dvb_index** out) {
    if (!ctx || !path || !*path || !out) { dvb_set_err(ctx, "dvb_chunk_index_load: invalid argument%s%s"); return DVB_E_INVALID; }
    *out = nullptr;
    dvb_scope sc(ctx);
    cudaStream_t st = sc.stream();
    FILE* f = fopen(path, "rb");
    if (!f) { dvb_set_err(ctx, "dvb_chunk_index_load: cannot open %s", path); return DVB_E_INVALID; }
    IxioHeader h;
    if (fread(&h, 1, sizeof(h), f) != sizeof(h) || memcmp(h.magic, IXIO_MAGIC, 8) != 0 || h.version != IXIO_VERSION || h.sub_bits != IXIO_SUB_BITS ||
        h.n_chunks < 0 || h.n_bytes < 0 || h.n_post < 0 || h.n_keys < 0 || h.n_segs == 0 || h.n_segs > (1u << 20)) {
        fclose(f); dvb_set_err(ctx, "dvb_chunk_index_load: %s is not a version-1 chunk index sidecar", path); return DVB_E_INVALID;
    }
    if (tag32 && memcmp(h.tag, tag32, 32) != 0) { fclose(f); dvb_set_err(ctx, "dvb_chunk_index_load: %s belongs to other content (tag mismatch)", path); return DVB_E_INVALID; }
    Bounce bb;
    if (!bb.ok()) { fclose(f); dvb_set_err(ctx, "dvb_chunk_index_load: cannot allocate the pinned bounce buffers%s%s"); return DVB_E_NOMEM; }
    dvb_index* ix = new dvb_index();
    ix->device = ctx->device; ix->n_chunks = h.n_chunks; ix->n_bytes = h.n_bytes; ix->n_post = h.n_post; ix->n_keys = h.n_keys;
    int32_t rc = DVB_OK; const char* why = "";
    int64_t next_base = 0, tot_post = 0, tot_keys = 0, tot_bytes = 0;
    for (uint32_t s = 0; s < h.n_segs && rc == DVB_OK; s++) {
        IxioSegHeader sh;
        if (fread(&sh, 1, sizeof(sh), f) != sizeof(sh)) { rc = DVB_E_INVALID; why = "truncated"; break; }
        if (sh.chunk_base != next_base || sh.n_chunks < 0 || sh.n_bytes < 0 || sh.n_post < 0 || sh.n_keys < 0 || sh.n_post > sh.n_bytes + 1 ||
            sh.n_keys > sh.n_post || sh.n_bytes > (1LL << 40) || sh.n_chunks > (1LL << 32)) { rc = DVB_E_INVALID; why = "inconsistent segment header"; break; }
        IxSeg sg; sg.chunk_base = sh.chunk_base; sg.n_chunks = sh.n_chunks; sg.n_bytes = sh.n_bytes; sg.n_post = sh.n_post; sg.n_keys = sh.n_keys;
        SegArrays a = seg_arrays(sg);
        void** slot[7] = { (void**)&sg.sym, (void**)&sg.sym_off, (void**)&sg.sym_cnt, (void**)&sg.pc, (void**)&sg.pv, (void**)&sg.ukeys, (void**)&sg.uoff };
        for (int i = 0; i < 7 && rc == DVB_OK; i++) {
            if (a.bytes[i] == 0 && i >= 3) continue;                                   // an index without postings has no sort arrays
            if (cudaMallocAsync(slot[i], a.bytes[i] + 64, st) != cudaSuccess) { cudaGetLastError(); rc = DVB_E_NOMEM; why = "out of device memory"; break; }
            if (!read_array(f, *slot[i], a.bytes[i], bb, st)) { rc = DVB_E_INVALID; why = "truncated or unreadable"; }
        }
        if (rc == DVB_OK) {
            uint64_t sum[7];
            if (!seg_checksums(ctx, sg, bb, st, sum)) { rc = DVB_E_CUDA; why = "checksum kernels failed"; }
            else for (int i = 0; i < 7; i++) if (sum[i] != sh.checksum[i]) { rc = DVB_E_INVALID; why = "checksum mismatch"; }
        }
        if (rc == DVB_OK && sg.n_chunks > 0) {
            int* d_bad = reinterpret_cast<int*>(bb.d_sum + 7); int h_bad = 0;
            cudaMemsetAsync(d_bad, 0, 4, st);
            k_ix_validate<<<ctx->num_sm * 8, 256, 0, st>>>(sg, d_bad);
            ctx->launches += 1;
            if (cudaMemcpyAsync(&h_bad, d_bad, 4, cudaMemcpyDeviceToHost, st) != cudaSuccess || cudaStreamSynchronize(st) != cudaSuccess) { rc = DVB_E_CUDA; why = "validation kernel failed"; }
            else if (h_bad) { rc = DVB_E_INVALID; why = "offsets or postings point outside the segment"; }
        }
        ix->segs.push_back(sg);                                                          // owned by ix from here on (freed below on error)
        next_base += sh.n_chunks; tot_post += sh.n_post; tot_keys += sh.n_keys; tot_bytes += sh.n_bytes;
    }
    fclose(f);
    if (rc == DVB_OK && (next_base != h.n_chunks || tot_post != h.n_post || tot_keys != h.n_keys || tot_bytes != h.n_bytes)) { rc = DVB_E_INVALID; why = "totals do not match the segments"; }
    cudaStreamSynchronize(st);
    if (rc != DVB_OK) {
        cudaGetLastError();
        for (IxSeg& sg : ix->segs) ix_free_seg(sg);
        delete ix;
        dvb_set_err(ctx, "dvb_chunk_index_load: %s: %s", path, why);
        return rc;
    }
    *out = ix;
    return DVB_OK;
}

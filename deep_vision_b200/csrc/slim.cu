// slim.cu — dvb_slim_facts: evidence selection for the report prompt, batched over reports.
// replaces: the scoring, ordering and three greedy passes of select_slimmed_facts_for_prompt
//           (reference web/server.py:22886-22993) for MANY evidence packs per launch; the host mirror
//           (deep_vision_b200/prompt_facts.py) keeps the dict handling, the reference-id sets and the signature strings.
//
// One warp per report.  Per fact the host passes what the reference computes from strings:
//   quality  = max(0.0, min(1.0, _safe_float(fact["quality_score"])))            (server.py:22923)
//   flags    bit 0 hard_triggered, bit 1 q_id in contradiction refs, bit 2 q_id in unknown refs, bit 3 is_follow_up
//   sig      dense id (0 .. n-1 within the report) of _fact_signature_for_prompt(fact)   (server.py:22876-22883)
//   dim      dense id (0 .. n-1 within the report) of the lower-cased dimension
//   key      _fact_sort_key_for_prompt(fact)                                      (server.py:22867-22874)
// Device: score = quality * 100 (+24 hard, +22 contradiction, +12 unknown, +8 follow-up, in that order, fp64 round to
// nearest per operation = Python's float arithmetic), order by (-score, index), the three passes of try_take with the
// reference's quota / quality / signature rules, then the stable sort of the selection by key.
#include "common.cuh"

#define SLIM_WARPS 4
#define SLIM_MAX_FACTS 1024              // facts per report held in shared memory (a report has ~10^2)

struct SlimWarp {
    double score[SLIM_MAX_FACTS];
    uint16_t order[SLIM_MAX_FACTS];      // rank -> fact
    uint16_t dimcnt[SLIM_MAX_FACTS];
    uint32_t sigtaken[SLIM_MAX_FACTS / 32];
};

__global__ void __launch_bounds__(SLIM_WARPS * 32) k_slim_facts(int64_t n_reports, const int64_t* __restrict__ fact_off, const double* __restrict__ quality,
                                                                 const uint8_t* __restrict__ flags, const int32_t* __restrict__ sig, const int32_t* __restrict__ dim,
                                                                 const int64_t* __restrict__ key, const int32_t* __restrict__ limit, const int64_t* __restrict__ out_off,
                                                                 int32_t dim_quota, double min_quality, int32_t keep_hard, int32_t dedup,
                                                                 int32_t* __restrict__ out_sel, int32_t* __restrict__ out_cnt) {
    extern __shared__ __align__(16) unsigned char slim_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    SlimWarp& w = reinterpret_cast<SlimWarp*>(slim_raw)[warp];
    for (int64_t r = (int64_t)blockIdx.x * SLIM_WARPS + warp; r < n_reports; r += (int64_t)gridDim.x * SLIM_WARPS) {
        const int64_t f0 = fact_off[r];
        const int n = (int)(fact_off[r + 1] - f0);
        const int lim = limit[r];
        int32_t* sel = out_sel + out_off[r];
        const int cap = (int)(out_off[r + 1] - out_off[r]);
        __syncwarp();
        // ---- scores (server.py:22930-22938)
        for (int i = lane; i < n; i += 32) {
            const uint32_t fl = flags[f0 + i];
            double s = __dmul_rn(quality[f0 + i], 100.0);
            if (fl & 1u) s = __dadd_rn(s, 24.0);
            if (fl & 2u) s = __dadd_rn(s, 22.0);
            if (fl & 4u) s = __dadd_rn(s, 12.0);
            if (fl & 8u) s = __dadd_rn(s, 8.0);
            w.score[i] = s;
            w.dimcnt[i] = 0;
        }
        for (int i = lane; i < (n + 31) / 32; i += 32) w.sigtaken[i] = 0u;
        __syncwarp();
        // ---- order by (-score, index): rank of fact i = facts that sort before it (server.py:22952)
        for (int i = lane; i < n; i += 32) {
            const double si = w.score[i];
            int rank = 0;
            for (int j = 0; j < n; j++) { const double sj = w.score[j]; rank += (sj > si || (sj == si && j < i)) ? 1 : 0; }
            w.order[rank] = (uint16_t)i;
        }
        __syncwarp();
        // ---- the three passes (server.py:22954-22989); serial by nature, lane 0 walks the order
        int cnt = 0;
        if (lane == 0) {
            for (int pass = 0; pass < 3 && cnt < lim; pass++) {
                const bool bypass = pass != 1;                            // passes 1 and 3 bypass quota and quality
                for (int k = 0; k < n; k++) {
                    if (cnt >= lim) break;
                    const int i = w.order[k];
                    const uint32_t fl = flags[f0 + i];
                    const bool mandatory = (fl & 2u) != 0u || (keep_hard && (fl & 1u) != 0u);
                    if (pass == 0 && !mandatory) continue;
                    const int sg = min(max(sig[f0 + i], 0), n - 1), dm = min(max(dim[f0 + i], 0), n - 1);     // ids are dense per report (documented); no fault on bad input
                    if (dedup && ((w.sigtaken[sg >> 5] >> (sg & 31)) & 1u)) continue;
                    if (!bypass && !mandatory && quality[f0 + i] < min_quality) continue;
                    if (!bypass && (int)w.dimcnt[dm] >= dim_quota) continue;
                    if (cnt < cap) sel[cnt] = i;
                    cnt++;
                    w.sigtaken[sg >> 5] |= 1u << (sg & 31);
                    w.dimcnt[dm] = (uint16_t)(w.dimcnt[dm] + 1);
                }
            }
        }
        cnt = __shfl_sync(0xffffffffu, cnt, 0);
        if (cnt > cap) cnt = cap;                                         // cannot happen with the host's capacity rule
        __syncwarp();
        // ---- stable sort of the selection by key (server.py:22992): rank among the selected, written through the score array
        int32_t* tmp = reinterpret_cast<int32_t*>(w.score);
        for (int a = lane; a < cnt; a += 32) {
            const int ia = sel[a]; const int64_t ka = key[f0 + ia];
            int rank = 0;
            for (int b = 0; b < cnt; b++) { const int64_t kb = key[f0 + sel[b]]; rank += (kb < ka || (kb == ka && b < a)) ? 1 : 0; }
            tmp[rank] = ia;
        }
        __syncwarp();
        for (int a = lane; a < cnt; a += 32) sel[a] = tmp[a];
        if (lane == 0) out_cnt[r] = cnt;
        __syncwarp();
    }
}

extern "C" int32_t dvb_slim_facts(dvb_ctx* ctx, int32_t mem, int64_t n_reports, const int64_t* fact_off, const double* quality, const uint8_t* flags,
                                   const int32_t* sig_id, const int32_t* dim_id, const int64_t* sort_key, const int32_t* limit, const int64_t* out_off,
                                   int32_t dim_quota, double min_quality, int32_t keep_hard, int32_t dedup, int32_t* out_sel, int32_t* out_cnt) {
    if (!ctx || n_reports < 0 || !fact_off || !limit || !out_off || !out_cnt || dim_quota < 1) {
        dvb_set_err(ctx, "dvb_slim_facts: invalid argument%s%s"); return DVB_E_INVALID; }
    if (n_reports == 0) return DVB_OK;
    dvb_scope sc(ctx); dvb_io io(sc, mem);
    std::vector<int64_t> fo((size_t)n_reports + 1), oo((size_t)n_reports + 1);
    if (mem == DVB_MEM_HOST) { memcpy(fo.data(), fact_off, 8 * (size_t)(n_reports + 1)); memcpy(oo.data(), out_off, 8 * (size_t)(n_reports + 1)); }
    else {
        DVB_CUDA(ctx, cudaMemcpyAsync(fo.data(), fact_off, 8 * (size_t)(n_reports + 1), cudaMemcpyDeviceToHost, io.st));
        DVB_CUDA(ctx, cudaMemcpyAsync(oo.data(), out_off, 8 * (size_t)(n_reports + 1), cudaMemcpyDeviceToHost, io.st));
        DVB_CUDA(ctx, cudaStreamSynchronize(io.st));
    }
    for (int64_t r = 0; r < n_reports; r++) {
        if (fo[r + 1] < fo[r] || oo[r + 1] < oo[r]) { dvb_set_err(ctx, "dvb_slim_facts: offsets must not decrease%s%s"); return DVB_E_INVALID; }
        if (fo[r + 1] - fo[r] > SLIM_MAX_FACTS || oo[r + 1] - oo[r] > 2 * SLIM_MAX_FACTS) {
            dvb_set_err(ctx, "dvb_slim_facts: a report has more than 1024 facts (or more than 2048 output slots)%s%s"); return DVB_E_CAPACITY; }
    }
    const int64_t n_facts = fo[n_reports] - fo[0], n_out = oo[n_reports] - oo[0];
    if (n_facts > 0 && (!quality || !flags || !sig_id || !dim_id || !sort_key)) { dvb_set_err(ctx, "dvb_slim_facts: null fact array%s%s"); return DVB_E_INVALID; }
    if (n_out > 0 && !out_sel) { dvb_set_err(ctx, "dvb_slim_facts: null output%s%s"); return DVB_E_INVALID; }
    // HOST: the fact arrays are staged from element fo[0] on, the device pointers are rebased so the offsets index them unchanged
    const int64_t* d_fo = io.in(fact_off, (size_t)n_reports + 1);
    const int64_t* d_oo = io.in(out_off, (size_t)n_reports + 1);
    const int32_t* d_lim = io.in(limit, (size_t)n_reports);
    const double* d_q = mem == DVB_MEM_HOST ? io.in(quality + fo[0], (size_t)n_facts) - fo[0] : quality;
    const uint8_t* d_fl = mem == DVB_MEM_HOST ? io.in(flags + fo[0], (size_t)n_facts) - fo[0] : flags;
    const int32_t* d_sg = mem == DVB_MEM_HOST ? io.in(sig_id + fo[0], (size_t)n_facts) - fo[0] : sig_id;
    const int32_t* d_dm = mem == DVB_MEM_HOST ? io.in(dim_id + fo[0], (size_t)n_facts) - fo[0] : dim_id;
    const int64_t* d_key = mem == DVB_MEM_HOST ? io.in(sort_key + fo[0], (size_t)n_facts) - fo[0] : sort_key;
    int32_t* d_sel = mem == DVB_MEM_HOST ? io.out(out_sel + oo[0], (size_t)n_out) - oo[0] : out_sel;
    int32_t* d_cnt = io.out(out_cnt, (size_t)n_reports);
    int32_t rc = io.check(ctx); if (rc) return rc;
    const size_t smem = sizeof(SlimWarp) * SLIM_WARPS;
    DVB_CUDA(ctx, cudaFuncSetAttribute(k_slim_facts, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int grid = (int)std::min<int64_t>((n_reports + SLIM_WARPS - 1) / SLIM_WARPS, (int64_t)ctx->num_sm * 4);
    k_slim_facts<<<grid, SLIM_WARPS * 32, smem, io.st>>>(n_reports, d_fo, d_q, d_fl, d_sg, d_dm, d_key, d_lim, d_oo, dim_quota, min_quality, keep_hard, dedup, d_sel, d_cnt);
    ctx->launches += 1;
    return io.finish(ctx);
}

"""Secondary configs of BASELINE.json measured on ONE B200 (per-GPU shard sizes where the config is
multi-GPU).  Writes one JSON line per measurement to gpurun_out/configs.jsonl.
    python scripts/bench_configs.py [c5] [small] [c3] [c4]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from omnia_b200 import GPUMemoryStore, FilterSpec, synth

PEAKS = json.load(open("MEASURED_PEAKS.json")) if os.path.exists("MEASURED_PEAKS.json") else {"hbm_gbs": 6650.0, "bf16_tflops_sustained": 1400.0, "bf16_tflops": 1590.0}
OUT = open("gpurun_out/configs.jsonl", "a")


def emit(d):
    print(json.dumps(d), flush=True)
    OUT.write(json.dumps(d) + "\n"); OUT.flush()


def timed_search(s, dq, k, steps=20, warm=3, **kw):
    nq = dq.shape[0]
    r = torch.empty((nq, k), dtype=torch.int64, device="cuda"); d = torch.empty((nq, k), dtype=torch.float64, device="cuda")
    c = torch.empty(nq, dtype=torch.int32, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    for _ in range(warm):
        s.device_search(dq, k, r, d, c, stream=stream, **kw)
    torch.cuda.synchronize()
    s.set_profiling(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        s.device_search(dq, k, r, d, c, stream=stream, **kw)
    e1.record(); torch.cuda.synchronize()
    st = s.stats(); s.set_profiling(False)
    return e0.elapsed_time(e1) / steps, st.scorer_ms_sum / max(1, st.scorer_timed), st


def c5():
    # C5: 50M x 384 over 8 GPUs, batch 1 -> one GPU's shard is 6.25M x 384 (4.8 GB bf16), HBM-bound
    n, dim = 6_250_000, 384
    corpus = synth.gauss_unit(n, dim)
    md = synth.metadata(n, n_workspaces=1000)
    with GPUMemoryStore(dim, n) as s:
        s.bulk_load(corpus, workspace_idx=md.workspace_idx, confidence=md.confidence, flags=md.flags,
                    valid_until_us=md.valid_until_us, workspaces=md.tables()["workspaces"])
        q = torch.from_numpy(synth.queries_from(corpus, 1)).cuda()
        words = s.bitmap_words()
        for name, flt in (("all live rows (no scope)", None), ("workspace ws-0000 (largest, Zipf)", FilterSpec("ws-0000", now_us=md.now_us))):
            bm = None
            if flt is not None:
                bm = torch.empty(words, dtype=torch.int32, device="cuda")
                s.device_build_bitmaps([flt], bm, torch.cuda.current_stream().cuda_stream)
            ms, sc, st = timed_search(s, q, 10, d_bitmaps=bm, n_groups=1 if bm is not None else 0)
            bytes_alg = n * dim * 2 + n / 8
            emit({"config": "C5 shard: 6.25M x 384 bf16, batch 1, top-10, " + name, "path": int(st.last_path),
                  "ms_per_query": ms, "queries_per_s": 1e3 / ms, "scorer_ms": sc,
                  "roofline": {"bound": "hbm", "achieved": bytes_alg / (sc / 1e3) / 1e9, "peak": PEAKS["hbm_gbs"], "unit": "GB/s",
                               "frac": bytes_alg / (sc / 1e3) / 1e9 / PEAKS["hbm_gbs"], "algorithmic_bytes": bytes_alg},
                  "fallback_queries": int(st.fallback_queries), "rescued_rows": int(st.rescued_rows)})


def small():
    n, dim = 1_000_000, 1536
    corpus = synth.gauss_unit(n, dim)
    with GPUMemoryStore(dim, n) as s:
        s.bulk_load(corpus)
        for nq in [int(x) for x in os.environ.get("BATCHES", "1,2,4,8,16,64,128,256,512,1024").split(",")]:
            q = torch.from_numpy(synth.queries_from(corpus, nq)).cuda()
            ms, sc, st = timed_search(s, q, 10)
            bytes_alg = n * dim * 2
            flops = 2.0 * nq * n * dim
            emit({"config": f"1M x 1536, batch {nq}, top-10", "path": int(st.last_path), "ms_per_batch": ms,
                  "queries_per_s": nq * 1e3 / ms, "scorer_ms": sc, "hbm_gbs": bytes_alg / (sc / 1e3) / 1e9,
                  "hbm_frac": bytes_alg / (sc / 1e3) / 1e9 / PEAKS["hbm_gbs"], "tflops": flops / (sc / 1e3) / 1e12,
                  "tensor_frac_sustained": flops / (sc / 1e3) / 1e12 / PEAKS["bf16_tflops_sustained"]})


def c3():
    # C3: 10M x 768 over 8 GPUs, top-100 -> one GPU's shard is 1.25M x 768
    n, dim, nq, k = 1_250_000, 768, 1024, 100
    corpus = synth.gauss_unit(n, dim)
    with GPUMemoryStore(dim, n) as s:
        s.bulk_load(corpus)
        q = torch.from_numpy(synth.queries_from(corpus, nq)).cuda()
        ms, sc, st = timed_search(s, q, k)
        flops = 2.0 * nq * n * dim
        emit({"config": "C3 shard: 1.25M x 768 bf16, 1024 queries, top-100 (local top-k of an 8-GPU 10M store)",
              "path": int(st.last_path), "sub_bucket": int(st.last_sub_bucket), "ms_per_batch": ms, "queries_per_s": nq * 1e3 / ms,
              "scorer_ms": sc, "roofline": {"bound": "tensor", "achieved": flops / (sc / 1e3) / 1e12, "peak": PEAKS["bf16_tflops_sustained"],
                                            "unit": "TFLOP/s", "frac": flops / (sc / 1e3) / 1e12 / PEAKS["bf16_tflops_sustained"]},
              "fallback_queries": int(st.fallback_queries), "rescued_rows": int(st.rescued_rows)})


def c4():
    # C4: near-duplicate self-join 1M x 1536, cosine >= 0.92
    n, dim = int(os.environ.get("C4_ROWS", 1_000_000)), 1536
    corpus, planted = synth.planted_dups(n, dim, n_pairs=2000)
    with GPUMemoryStore(dim, n, force_path=int(os.environ.get("C4_PATH", 0))) as s:
        s.bulk_load(corpus)
        from bench import ClockSampler
        for sbv in os.environ.get("C4_SB", "default").split(","):
          if sbv != "default": os.environ["OMR_SELFJOIN_SB"] = sbv
          for rep in range(2):
              s.set_profiling(True)
              st0 = s.stats()
              cs = ClockSampler(0, 0.02); cs.start()
              t0 = time.perf_counter()
              pairs = s.run_entity_duplicate_candidates("default", 0.92, 0)
              dt = time.perf_counter() - t0
              clocks = cs.stop()
              kernel_ms = s.stats().scorer_ms_sum - st0.scorer_ms_sum
              s.set_profiling(False)
          want = {(min(int(i), int(j)), max(int(i), int(j))) for i, j, sim in planted if sim >= float(np.float32(0.92))}
          got = {(min(p.canonical_row, p.dupe_row), max(p.canonical_row, p.dupe_row)) for p in pairs}
          flops = 2.0 * dim * n * (n + 256) / 2
          emit({"config": f"C4: self-join {n} x 1536, cosine >= 0.92", "seconds": dt, "pairs": len(pairs),
                "kernel_ms": kernel_ms, "clocks": clocks, "sub_block": os.environ.get("OMR_SELFJOIN_SB", "default"), "kernel": "1 cta" if os.environ.get("C4_PATH") == "2" else "cta pair", "planted_above_floor": len(want), "planted_found": len(want & got), "unplanted_pairs": len(got - want),
                "roofline": {"bound": "tensor", "achieved": flops / dt / 1e12, "peak": PEAKS["bf16_tflops_sustained"], "unit": "TFLOP/s",
                             "frac": flops / dt / 1e12 / PEAKS["bf16_tflops_sustained"], "note": "wall time of the whole omr_selfjoin call, host post-processing included"}})


def pca():
    # section 8(f)-4: the Memory Galaxy PCA input.  Default galaxy = 2000 entities x 1536 -> 50 columns;
    # the large case shows where the kernel, not the 19 MB covariance read-back, sets the time.
    dim = 1536
    for n in (2000, 200_000):
        x = synth.clustered(n, dim, seed=1)
        with GPUMemoryStore(dim, n) as s:
            s.bulk_load(x)
            rows = np.arange(n, dtype=np.int64)
            for rep in range(3):
                t0 = time.perf_counter(); mean, cov = s.pca_moments(rows); t_mom = time.perf_counter() - t0
            t0 = time.perf_counter(); w, v = np.linalg.eigh(cov); t_eig = time.perf_counter() - t0
            basis = np.ascontiguousarray(v[:, np.argsort(-w, kind="stable")[:50]])
            for rep in range(3):
                t0 = time.perf_counter(); out = s.pca_project(rows, basis); t_proj = time.perf_counter() - t0
        x64 = x.astype(np.float64)
        t0 = time.perf_counter(); c = x64 - x64.mean(axis=0); ref_cov = c.T @ c / (n - 1); t_cpu_cov = time.perf_counter() - t0
        err = float(np.abs(cov - ref_cov).max())
        t_svd = None
        if n <= 2000:
            t0 = time.perf_counter(); np.linalg.svd(c, full_matrices=False); t_svd = time.perf_counter() - t0
        flops = 2.0 * n * dim * (dim + 128) / 2          # upper-triangle tiles of 128
        emit({"config": f"PCA moments {n} x {dim} (fp64)", "moments_s": t_mom, "host_eigh_s": t_eig, "project_s": t_proj,
              "cpu_numpy_cov_s": t_cpu_cov, "cpu_numpy_svd_s": t_svd, "max_abs_cov_err": err, "cpu_threads": os.cpu_count(),
              "roofline": {"bound": "fp64", "achieved": flops / t_mom / 1e12, "peak": 40.0, "unit": "TFLOP/s",
                           "frac": flops / t_mom / 1e12 / 40.0,
                           "note": "wall time of omr_pca_moments incl. the covariance D2H; peak is the nominal B200 fp64 figure, unmeasured"}})


if __name__ == "__main__":
    os.makedirs("gpurun_out", exist_ok=True)
    which = sys.argv[1:] or ["c5", "small", "c3", "c4"]
    for w in which:
        {"c5": c5, "small": small, "c3": c3, "c4": c4, "pca": pca}[w]()

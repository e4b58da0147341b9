"""End-to-end check on the reference's own pipeline (BASELINE.json configs[0], SURVEY.md 8d "G1"):
synthetic chr20 FASTA + 30x BAM -> `--analysis getCIGARindels` -> windows -> `--analysis indels
--doDiploid` -> `.glf.txt` -> VCF.

  * CPU: the window scripts' restatements (oracle/e2e/make_windows.py, merge_diploid.py), and the stock
    reference binary (oracle/_ref/dindel_ref, built where /root/reference exists) recovering the planted
    indels from the synthetic BAM -- this pins the fixture generator and the harness.
  * GPU: the reference binary with INTEGRATION.md's patch (oracle/_ref/dindel_b200: K1, K4 and the inference cores
    K2 / K3 run on the B200 through include/dindel_b200.hpp) produces `.glf.txt` files and a VCF that are byte-identical
    to the stock CPU binary's (north star: "final VCF indel calls and genotypes must be bit-exact").
"""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
E2E = os.path.join(ROOT, "oracle", "e2e")
sys.path.insert(0, E2E)
import make_windows  # noqa: E402
import merge_diploid  # noqa: E402
import run_e2e  # noqa: E402


def test_select_and_make_windows(tmp_path):
    src = tmp_path / "v.txt"
    src.write_text("20 100 +A # 1\n20 100 +A -CG # 2 3\n20 105 -T # 2\n20 400 +GG # 1\n20 900 -AAAA # 5\n")
    sel = tmp_path / "sel.txt"
    make_windows.select_candidates(str(src), str(sel), 2)
    assert sel.read_text() == "20 100 +A # 3\n20 100 -CG # 3\n20 105 -T # 2\n20 900 -AAAA # 5\n"
    files = make_windows.split_and_merge(str(sel), str(tmp_path / "w"), 20, 1000)
    assert len(files) == 1
    lines = open(files[0]).read().splitlines()
    # 100 and 105 cluster (<= minDist apart); window = [min refStart - 60, max refEnd + 60]
    assert lines[0] == "20 40 165 100,+A 100,-CG 105,-T"
    assert lines[1] == "20 840 963 900,-AAAA"


def test_merge_restatement_vcf_line(tmp_path):
    fa = tmp_path / "r.fa"
    seq = "ACGTTGCA" * 20
    fa.write_text(">20\n" + "\n".join(seq[i:i + 60] for i in range(0, len(seq), 60)) + "\n")
    (tmp_path / "r.fa.fai").write_text("20\t%d\t4\t60\t61\n" % len(seq))
    hdr = ("msg index analysis_type tid lpos rpos center_position realigned_position was_candidate_in_window "
           "ref_all nref_all num_reads post_prob_variant qual est_freq logZ hapfreqs indidx msq numOffAll num_indel "
           "num_cover_forward num_cover_reverse num_unmapped_realigned var_coverage_forward var_coverage_reverse "
           "nBQT nmmBQT mLogBQ nMMLeft nMMRight glf")
    row = ("ok 1 dip.map 20 10 130 70 64 1 NA -GT 50 NA 321.9 NA NA NA 0 57.1 NA NA 18 24 0 8 12 NA NA NA NA NA "
           "0/1:143.7")
    glf = tmp_path / "a.glf.txt"
    glf.write_text(hdr + "\n" + row + "\n")
    n = merge_diploid.merge_output([str(glf)], str(fa), str(tmp_path / "o.vcf"))
    assert n == 1
    last = open(tmp_path / "o.vcf").read().splitlines()[-1].split("\t")
    # realigned_position 64 (0-based) -> fetch 1-based pos 64 = seq[63:66]; deletion of the 2 bases after it
    assert last[0:2] == ["20", "64"] and last[3] == seq[63:66] and last[4] == seq[63]
    assert last[5] == "321" and last[6] == "PASS" and last[-1] == "0/1:143"
    assert last[7].startswith("DP=50;NF=8;NR=12;NRS=18;NFS=24;HP=")


def _truth_indels(path):
    out = []
    for line in open(path):
        chrom, pos, kind, ln, alt, gt = line.split()
        if kind in ("ins", "del"):
            out.append((int(pos), kind, int(ln), gt))
    return out


def test_cpu_reference_pipeline_recovers_planted_indels(tmp_path):
    if not run_e2e.have_binaries(("cpu",)):
        pytest.skip("oracle/_ref/dindel_ref not built (reference sources absent)")
    work = str(tmp_path)
    prep = run_e2e.prepare(work, 40000, seed=3, per_file=8)
    assert prep["n_windows"] >= 10 and len(prep["window_files"]) >= 2
    glfs, vcf, wall, n_calls = run_e2e.run_arm("cpu", prep, work, procs=4)
    calls = {}
    for line in open(vcf):
        if line.startswith("#"):
            continue
        d = line.split("\t")
        if len(d[3]) != len(d[4].split(",")[0]):
            calls[int(d[1])] = (d[3], d[4], d[9].split(":")[0])
    truth = _truth_indels(os.path.join(work, "truth.txt"))
    assert len(truth) >= 10
    hit = 0
    for pos, kind, ln, gt in truth:
        for p in range(pos - 12, pos + 3):   # dindel left-aligns; VCF pos is the base before the event
            if p in calls:
                ref, alt, g = calls[p]
                if (len(ref) - len(alt) == ln if kind == "del" else len(alt) - len(ref) == ln):
                    hit += (g == ("1/1" if gt == "1|1" else "0/1"))
                    break
    assert hit >= 0.8 * len(truth), (hit, len(truth))


def test_patched_binary_has_no_cpu_path(tmp_path):
    """Without a CUDA device every window of the GPU-backed binary is skipped with the library's error;
    nothing falls back to the reference's CPU loop."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA device present")
    if not run_e2e.have_binaries(("cpu", "gpu")):
        pytest.skip("oracle/_ref binaries not built")
    work = str(tmp_path)
    prep = run_e2e.prepare(work, 20000, seed=5)
    glfs, vcf, wall, n_calls = run_e2e.run_arm("gpu", prep, work, procs=1)
    assert n_calls == 0
    body = [l for l in open(glfs[0]).read().splitlines()[1:]]
    assert body and all(("no_usable_CUDA_device" in l) or l.startswith("error_too_few_reads") for l in body), body[:3]


@pytest.mark.gpu
def test_glf_and_vcf_identical_to_cpu_reference(tmp_path):
    if not run_e2e.have_binaries(("cpu", "gpu")):
        pytest.skip("oracle/_ref/dindel_ref / dindel_b200 not built (reference sources absent at build time)")
    r = subprocess.run([sys.executable, os.path.join(E2E, "run_e2e.py"), "--work", str(tmp_path), "--length", "120000",
                        "--seed", "2", "--arms", "cpu,gpu,check", "--windows-per-file", "20", "--heap-neutral"],
                       capture_output=True, text=True, timeout=1500)
    assert r.returncode == 0, r.stdout + r.stderr
    res = json.loads(r.stdout.strip().splitlines()[-1])
    # every pair the real pipeline produced: GPU MLAlignment == reference class's, field by field
    chk = res["default"]["check"]
    assert chk["pair_mismatches"] == 0 and chk["pairs_compared"] >= 10000, res
    # glf + VCF byte-identical with the reference's stray hpos[size()] read neutralised in both arms
    hn = res["heap_neutral"]
    assert hn["glf_identical"] and hn["vcf_identical"], res
    assert hn["cpu"]["vcf_calls"] >= 20 and hn["cpu"]["windows_ok"] >= 40, res
    # diploidGLF's read sums came from K2 (dindel_b200_process); the guard band sent few windows back to the CPU loops
    inf = hn["gpu"]["inference"]
    assert inf["windows"] >= 40 and inf["rechecked_on_host"] <= inf["windows"] // 2, inf
    print("K2 inference:", inf)


@pytest.mark.gpu
def test_glf_identical_with_filter_haplotypes(tmp_path):
    """--filterHaplotypes: filtered[] from K4 (dindel_b200_coverage) now changes which haplotype pairs diploidGLF
    considers; output still byte-identical to the stock CPU binary."""
    if not run_e2e.have_binaries(("cpu", "gpu")):
        pytest.skip("oracle/_ref/dindel_ref / dindel_b200 not built (reference sources absent at build time)")
    r = subprocess.run([sys.executable, os.path.join(E2E, "run_e2e.py"), "--work", str(tmp_path), "--length", "60000",
                        "--seed", "4", "--arms", "cpu,gpu", "--windows-per-file", "20", "--heap-neutral",
                        "--filter-haplotypes"], capture_output=True, text=True, timeout=1500)
    assert r.returncode == 0, r.stdout + r.stderr
    res = json.loads(r.stdout.strip().splitlines()[-1])
    hn = res["heap_neutral"]
    assert hn["glf_identical"] and hn["vcf_identical"] and hn["cpu"]["windows_ok"] >= 20, res


@pytest.mark.gpu
def test_glf_identical_with_lib_file(tmp_path):
    """--libFile (the README pipeline): insert-size prior in computeBMidPrior (ObservationModelFB.cpp:279-293) for every
    paired read + mate rescue of unmapped reads on the host; the GPU arm gets the libraries through
    dindel_b200_set_libraries.  Every pipeline pair equals the reference class's, glf and VCF byte-identical."""
    if not run_e2e.have_binaries(("cpu", "gpu", "check")):
        pytest.skip("oracle/_ref binaries not built (reference sources absent at build time)")
    r = subprocess.run([sys.executable, os.path.join(E2E, "run_e2e.py"), "--work", str(tmp_path), "--length", "60000",
                        "--seed", "6", "--arms", "cpu,gpu,check", "--windows-per-file", "20", "--heap-neutral",
                        "--lib-file"], capture_output=True, text=True, timeout=1500)
    assert r.returncode == 0, r.stdout + r.stderr
    res = json.loads(r.stdout.strip().splitlines()[-1])
    chk = res["default"]["check"]
    assert chk["pair_mismatches"] == 0 and chk["pairs_compared"] >= 5000, res
    hn = res["heap_neutral"]
    assert hn["glf_identical"] and hn["vcf_identical"] and hn["cpu"]["windows_ok"] >= 20, res


@pytest.mark.gpu
def test_glf_identical_pooled(tmp_path):
    """--doPooled (estimateHaplotypeFrequenciesBayesEM, the path of configs[2]): K1, K4, the EM (K3) and the pooled
    genotype sums (K2) on the GPU; glf byte-identical to the stock CPU binary (both arms use the same digamma shim)."""
    if not run_e2e.have_binaries(("cpu", "gpu")):
        pytest.skip("oracle/_ref binaries not built (reference sources absent at build time)")
    r = subprocess.run([sys.executable, os.path.join(E2E, "run_e2e.py"), "--work", str(tmp_path), "--length", "50000",
                        "--seed", "8", "--arms", "cpu,gpu", "--windows-per-file", "20", "--heap-neutral", "--pooled"],
                       capture_output=True, text=True, timeout=1500)
    assert r.returncode == 0, r.stdout + r.stderr
    res = json.loads(r.stdout.strip().splitlines()[-1])
    hn = res["heap_neutral"]
    assert hn["glf_identical"] and hn["cpu"]["windows_ok"] >= 15, res
    # the EM ran on K3 and the pooled genotype sums on K2; knife-edge sets / fields went back to the CPU loops
    inf = hn["gpu"]["inference"]
    assert inf["windows"] >= 15 and inf["em_sets"] >= 30 and inf["em_set_mismatch"] == 0, inf
    assert inf["rechecked_on_host"] <= inf["windows"] // 2, inf
    print("K3 inference:", inf)


@pytest.mark.gpu
@pytest.mark.parametrize("pooled", [False, True])
def test_realigned_bams_identical(tmp_path, pooled):
    """--outputRealignedBAM: the per-window realigned BAMs of the GPU-backed binary (CIGARs and positions from K5,
    dindel_b200_cigars, the fourth patch line) equal the stock CPU binary's record by record, --doDiploid (haplotype
    choice among the MAP pair, DInDel.cpp:597-606) and --doPooled (best haplotype, :503-509); glf identical too."""
    if not run_e2e.have_binaries(("cpu", "gpu")):
        pytest.skip("oracle/_ref binaries not built (reference sources absent at build time)")
    r = subprocess.run([sys.executable, os.path.join(E2E, "run_e2e.py"), "--work", str(tmp_path), "--length", "40000",
                        "--seed", "12", "--arms", "cpu,gpu", "--windows-per-file", "20", "--realigned-bam"] +
                       (["--pooled"] if pooled else []), capture_output=True, text=True, timeout=1500)
    assert r.returncode == 0, r.stdout + r.stderr
    res = json.loads(r.stdout.strip().splitlines()[-1])
    d = res["default"]
    rb = d["realigned_bam"]
    assert rb["mismatches"] == 0 and rb["files"] >= 10 and rb["records"] >= 500 and rb["records_with_indel_or_clip"] >= 50, res
    assert d["glf_identical"], res


@pytest.mark.gpu
@pytest.mark.parametrize("extra", [[], ["--filter-haplotypes"], ["--lib-file"], ["--pooled"]])
def test_batched_driver_one_process_identical(tmp_path, extra):
    """The batched window loop (include/dindel_b200_driver.hpp): ONE GPU-backed process over every window -- getReads /
    getHaplotypes on host threads, one dindel_b200_process (K1 + K4 + K2) per batch of windows, the reference's own
    empiricalDistributionMethod replayed per window on host threads -- writes the glf lines of the stock binary run one
    process per window file (the `index` column, a window's line number inside its file, aside) and the same VCF.  Small
    batches (DINDEL_B200_BATCH) force several flushes."""
    if not run_e2e.have_binaries(("cpu", "gpu")):
        pytest.skip("oracle/_ref binaries not built (reference sources absent at build time)")
    env = dict(os.environ, DINDEL_B200_BATCH="23")
    r = subprocess.run([sys.executable, os.path.join(E2E, "run_e2e.py"), "--work", str(tmp_path), "--length", "80000",
                        "--seed", "21", "--arms", "cpu,gpu", "--windows-per-file", "15", "--heap-neutral", "--gpu-one-file",
                        "--gpu-threads", "5"] + extra, capture_output=True, text=True, timeout=1500, env=env)
    assert r.returncode == 0, r.stdout + r.stderr
    res = json.loads(r.stdout.strip().splitlines()[-1])
    hn = res["heap_neutral"]
    assert hn["glf_identical"] and hn["cpu"]["windows_ok"] >= 25, res
    if "--pooled" not in extra:
        assert hn["vcf_identical"] and hn["cpu"]["vcf_calls"] >= 15, res
    lines = [l for l in hn["gpu"]["batched"] if l.startswith("dindel_b200 batched:")]
    assert lines and any(l.startswith("dindel_b200 host stages") for l in hn["gpu"]["batched"]), res
    nb = int(lines[0].split("batches")[1].split()[0])
    assert nb >= 2, hn["gpu"]["batched"]
    assert hn["gpu"]["inference"]["windows"] >= 25, res


@pytest.mark.gpu
@pytest.mark.parametrize("pooled", [False, True])
def test_window_at_a_time_loop_still_identical(tmp_path, pooled):
    """DINDEL_B200_NO_BATCH: the reference's own window loop with one GPU call per window,
    K2 / K3 through Engine::computeLikelihoods / Engine::windowEM."""
    if not run_e2e.have_binaries(("cpu", "gpu")):
        pytest.skip("oracle/_ref binaries not built (reference sources absent at build time)")
    r = subprocess.run([sys.executable, os.path.join(E2E, "run_e2e.py"), "--work", str(tmp_path), "--length", "40000",
                        "--seed", "23", "--arms", "cpu,gpu", "--windows-per-file", "20", "--heap-neutral", "--gpu-no-batch"] +
                       (["--pooled"] if pooled else []), capture_output=True, text=True, timeout=1500)
    assert r.returncode == 0, r.stdout + r.stderr
    res = json.loads(r.stdout.strip().splitlines()[-1])
    hn = res["heap_neutral"]
    assert hn["glf_identical"] and hn["cpu"]["windows_ok"] >= 12, res
    assert hn["gpu"]["inference"]["windows"] >= 12 and not hn["gpu"]["batched"], res


@pytest.mark.gpu
@pytest.mark.parametrize("extra", [["--gpu-one-file"], ["--gpu-one-file", "--pooled"], ["--gpu-no-batch"], ["--gpu-no-batch", "--pooled"]])
def test_device_read_stats_identical(tmp_path, extra):
    """K6 in the pipeline (DINDEL_B200_DEVICE_STATS=1: every site, whatever its size): the read statistics behind every
    called site -- dip.map sites, variant positions, pooled variants (DInDel.cpp:3126-3170, :3396-3404, :3504-3554,
    :2697-2742) -- come from dindel_b200_read_stats instead of the reference's loops; the glf columns they feed (msq,
    num_cover_forward / reverse, nBQT, nmmBQT, mLogBQ, nMMLeft, nMMRight, numOffAll, num_indel, num_unmapped_realigned)
    and the VCF stay byte-identical, in the batched driver and in the window-at-a-time loop."""
    if not run_e2e.have_binaries(("cpu", "gpu")):
        pytest.skip("oracle/_ref binaries not built (reference sources absent at build time)")
    env = dict(os.environ, DINDEL_B200_BATCH="30")
    r = subprocess.run([sys.executable, os.path.join(E2E, "run_e2e.py"), "--work", str(tmp_path), "--length", "60000",
                        "--seed", "41", "--arms", "cpu,gpu", "--windows-per-file", "15", "--heap-neutral", "--gpu-threads", "4",
                        "--gpu-device-stats", "1"] + extra, capture_output=True, text=True, timeout=1500, env=env)
    assert r.returncode == 0, r.stdout + r.stderr
    res = json.loads(r.stdout.strip().splitlines()[-1])
    hn = res["heap_neutral"]
    assert hn["glf_identical"] and hn["cpu"]["windows_ok"] >= 18, res
    if "--pooled" not in extra:
        assert hn["vcf_identical"] and hn["cpu"]["vcf_calls"] >= 10, res
    rs = hn["gpu"]["read_stats"]
    assert rs["device_calls"] >= 18 and rs["jobs"] >= rs["device_calls"], res


@pytest.mark.gpu
@pytest.mark.parametrize("pooled", [False, True])
def test_long_reads_pipeline_identical(tmp_path, pooled):
    """configs[3] through the real binary: 250-bp reads (longer than most candidate haplotypes: long LO / RO overhangs),
    --maxLengthIndel 10 (the J = 11 kernel classes) and up to 32 haplotypes per window; glf (and VCF) byte-identical to the
    stock CPU binary."""
    if not run_e2e.have_binaries(("cpu", "gpu")):
        pytest.skip("oracle/_ref binaries not built (reference sources absent at build time)")
    r = subprocess.run([sys.executable, os.path.join(E2E, "run_e2e.py"), "--work", str(tmp_path), "--length", "30000",
                        "--coverage", "40", "--seed", "9", "--read-len", "250", "--stage3-args", "--maxLengthIndel 10 --maxHap 32",
                        "--arms", "cpu,gpu", "--windows-per-file", "3", "--heap-neutral", "--gpu-one-file"] +
                       (["--pooled"] if pooled else []), capture_output=True, text=True, timeout=1500)
    assert r.returncode == 0, r.stdout + r.stderr
    res = json.loads(r.stdout.strip().splitlines()[-1])
    hn = res["heap_neutral"]
    assert hn["glf_identical"] and hn["cpu"]["windows_ok"] >= 10 and hn["gpu"]["inference"]["windows"] >= 10, res
    if not pooled:
        assert hn["vcf_identical"] and hn["cpu"]["vcf_calls"] >= 8, res


@pytest.mark.gpu
def test_batched_driver_over_several_devices(tmp_path):
    """configs[4] in the small: the patched binary's batched driver shards every batch of windows over several CUDA
    devices through the in-process scheduler (DINDEL_B200_DEVICES -> Engine::enableMulti -> dindel_b200_multi_process);
    glf and VCF byte-identical to the stock CPU binary.  Two contexts on device 0 where there is one GPU, every visible
    device (up to 8) otherwise."""
    import torch
    if not run_e2e.have_binaries(("cpu", "gpu")):
        pytest.skip("oracle/_ref binaries not built (reference sources absent at build time)")
    n = min(torch.cuda.device_count(), 8)
    devs = ",".join(str(i) for i in range(n)) if n >= 2 else "0,0"
    env = dict(os.environ, DINDEL_B200_BATCH="40")
    r = subprocess.run([sys.executable, os.path.join(E2E, "run_e2e.py"), "--work", str(tmp_path), "--length", "150000",
                        "--seed", "31", "--arms", "cpu,gpu", "--windows-per-file", "25", "--heap-neutral", "--gpu-one-file",
                        "--gpu-devices", devs], capture_output=True, text=True, timeout=1500, env=env)
    assert r.returncode == 0, r.stdout + r.stderr
    res = json.loads(r.stdout.strip().splitlines()[-1])
    hn = res["heap_neutral"]
    assert hn["glf_identical"] and hn["vcf_identical"] and hn["cpu"]["windows_ok"] >= 50 and hn["cpu"]["vcf_calls"] >= 30, res
    assert hn["gpu"]["inference"]["windows"] >= 50, res

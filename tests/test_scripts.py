"""Shell / deployment layer: syntax, placeholder substitution, env -> argv mapping of
the container entrypoint, result extraction.  CPU only; kubectl, docker and GPUs are never touched."""
import glob
import json
import os
import re
import stat
import subprocess

import pytest
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHELL_FILES = sorted(glob.glob(os.path.join(ROOT, "scripts", "*.sh"))) + [
    os.path.join(ROOT, "docker", "entrypoint.sh"), os.path.join(ROOT, "baseline", "install_reference.sh")]


@pytest.mark.parametrize("path", SHELL_FILES, ids=[os.path.basename(p) for p in SHELL_FILES])
def test_shell_syntax_and_executable_bit(path):
    assert subprocess.run(["bash", "-n", path], capture_output=True, text=True).returncode == 0
    assert os.stat(path).st_mode & stat.S_IXUSR, f"{path} is not executable"


def test_reference_script_set_is_present():
    # every script of the reference's scripts/ directory has a counterpart
    names = {os.path.basename(p) for p in glob.glob(os.path.join(ROOT, "scripts", "*"))}
    for n in ("run_all_benchmarks.sh", "launch_multi.sh", "launch_smoke.sh", "collect_results.sh",
              "check_cluster_gpus.sh", "verify_offline.sh", "build.sh", "push.sh", "install_analysis_deps.sh",
              "parse_metrics.py", "plot.py", "make_report.py"):
        assert n in names, n


SUBST = {"JOB": "bench-ddp-ws8-seq2048-tierA", "STRATEGY": "ddp", "WORLD_SIZE": "8", "WORKER_COUNT": "7",
         "SEQ_LEN": "2048", "TIER": "A", "STEPS": "100", "PER_DEVICE_BATCH": "1", "GRAD_ACCUM": "4",
         "IMAGE": "registry.example/b200bench:test"}


@pytest.mark.parametrize("name", ["job-master.template.yaml", "job-workers.template.yaml", "job-multi.template.yaml",
                                  "job-smoke-1gpu.yaml", "job-single.tmpl.yaml", "namespace.yaml",
                                  "serviceaccount.yaml", "service-master.yaml"])
def test_k8s_manifests_render_to_valid_yaml(name):
    text = open(os.path.join(ROOT, "k8s", name)).read()
    for k, v in SUBST.items():
        text = text.replace("{{%s}}" % k, v)
    text = text.replace("__IMAGE__", SUBST["IMAGE"])
    assert not re.search(r"\{\{[A-Z_]+\}\}", text), "unsubstituted placeholder left"
    docs = [d for d in yaml.safe_load_all(text) if d]
    assert docs and all("kind" in d and "apiVersion" in d for d in docs)
    for d in docs:
        if d["kind"] == "Job":
            pod = d["spec"]["template"]["spec"]
            assert pod["restartPolicy"] == "Never"
            c = pod["containers"][0]
            assert "nvidia.com/gpu" in c["resources"]["limits"]


def test_launch_multi_placeholders_match_template():
    sh = open(os.path.join(ROOT, "scripts", "launch_multi.sh")).read()
    tmpl = open(os.path.join(ROOT, "k8s", "job-multi.template.yaml")).read()
    assert set(re.findall(r"\{\{([A-Z_]+)\}\}", tmpl)) <= set(re.findall(r"\{\{([A-Z_]+)\}\}", sh))
    r = subprocess.run(["bash", os.path.join(ROOT, "scripts", "launch_multi.sh"), "--world-size", "1"],
                       capture_output=True, text=True)
    assert r.returncode == 1 and "world-size" in r.stderr                 # reference: ws >= 2 required
    r = subprocess.run(["bash", os.path.join(ROOT, "scripts", "launch_multi.sh"), "--bogus"],
                       capture_output=True, text=True)
    assert r.returncode == 1


def _run_entrypoint(tmp_path, env_extra):
    """Runs docker/entrypoint.sh with stub `python` and `nvidia-smi` first in PATH; returns the argv python got."""
    bindir = tmp_path / "bin"
    bindir.mkdir(exist_ok=True)
    out = tmp_path / "argv.json"
    (bindir / "python").write_text(
        "#!/bin/bash\n/usr/bin/env python3 -c 'import json,sys; json.dump(sys.argv[1:], open(sys.argv[1], \"w\"))' "
        f"{out} \"$@\"\n")
    (bindir / "nvidia-smi").write_text("#!/bin/bash\necho 'GPU 0: stub'\n")
    for f in ("python", "nvidia-smi"):
        os.chmod(bindir / f, 0o755)
    env = {"PATH": f"{bindir}:{os.environ['PATH']}", "APP_DIR": ROOT, "RESULTS_DIR": str(tmp_path / "res")}
    env.update(env_extra)
    r = subprocess.run(["bash", os.path.join(ROOT, "docker", "entrypoint.sh")], env=env, capture_output=True,
                       text=True, timeout=60)
    assert r.returncode == 0, r.stderr
    return json.load(open(out))[1:]


def _flag(argv, name):
    return argv[argv.index(name) + 1]


def test_entrypoint_rank_per_pod_env_contract(tmp_path):
    # indexed worker job: RANK = JOB_COMPLETION_INDEX + 1 (reference entrypoint.sh:11-16)
    argv = _run_entrypoint(tmp_path, {"STRATEGY": "zero3", "WORLD_SIZE": "4", "JOB_COMPLETION_INDEX": "2",
                                      "MASTER_ADDR": "bench-master.bench.svc.cluster.local", "GPUS_PER_POD": "1",
                                      "SEQ_LEN": "1024", "TIER": "B", "STEPS": "7", "GRAD_ACCUM": "4"})
    assert argv[1].endswith("benchmarking/train_harness.py")
    assert _flag(argv, "--rank") == "3" and _flag(argv, "--world-size") == "4"
    assert _flag(argv, "--master-addr") == "bench-master.bench.svc.cluster.local"
    assert _flag(argv, "--deepspeed-config").endswith("configs/deepspeed/zero3.json")
    assert (_flag(argv, "--seq-len"), _flag(argv, "--tier"), _flag(argv, "--steps")) == ("1024", "B", "7")
    assert "--synthetic" in argv
    # rank 0 advertises its own pod IP (reference entrypoint.sh:21-26)
    argv = _run_entrypoint(tmp_path, {"STRATEGY": "fsdp", "WORLD_SIZE": "2", "RANK": "0", "POD_IP": "10.1.2.3",
                                      "MASTER_ADDR": "ignored", "GPUS_PER_POD": "1"})
    assert _flag(argv, "--master-addr") == "10.1.2.3" and _flag(argv, "--rank") == "0"
    assert _flag(argv, "--fsdp-config").endswith("configs/fsdp/fsdp_config.yaml")


def test_entrypoint_single_host_mode_uses_the_local_launcher(tmp_path):
    argv = _run_entrypoint(tmp_path, {"STRATEGY": "ddp", "WORLD_SIZE": "8", "GPUS_PER_POD": "8"})
    assert argv[:3] == ["-u", "-m", "b200bench.launch"] and _flag(argv, "--nproc") == "8"
    assert "--deepspeed-config" not in argv and "--fsdp-config" not in argv


def test_collect_results_cuts_the_marked_json(tmp_path):
    job = "bench-ddp-ws1-seq2048-tierA"
    payload = {"strategy": "ddp", "world_size": 1, "tokens_per_sec": 123.5}
    (tmp_path / f"{job}.log").write_text(
        "noise\nBENCHMARK_RESULT_JSON_START\n" + json.dumps(payload, indent=2) + "\nBENCHMARK_RESULT_JSON_END\nbye\n")
    r = subprocess.run(["bash", os.path.join(ROOT, "scripts", "collect_results.sh"), "local", job, str(tmp_path)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert json.load(open(tmp_path / f"{job}_results" / "result.json")) == payload
    # a log without markers is an error (the reference's suite counts it as completed)
    (tmp_path / "bad.log").write_text("no markers here\n")
    r = subprocess.run(["bash", os.path.join(ROOT, "scripts", "collect_results.sh"), "local", "bad", str(tmp_path)],
                       capture_output=True, text=True)
    assert r.returncode != 0


def test_coverage_document_points_at_existing_files():
    """docs/COVERAGE.md maps every inventory row of SURVEY.md to files and tests: none of them may have moved away."""
    import re
    text = open(os.path.join(ROOT, "docs", "COVERAGE.md")).read()
    prefixes = ["", "b200bench/", "b200bench/parallel/", "b200bench/utils/", "b200bench/models/", "csrc/", "scripts/",
                "tests/", "tests/multi/", "k8s/", "profiles/", "docs/"]
    missing = []
    for m in sorted(set(re.findall(r"`([A-Za-z0-9_./\-]+\.(?:py|cu|cuh|h|sh|md|json|yaml|log))`", text))):
        if m.startswith("zero2.json"):
            continue                                  # citation of a reference file with line numbers
        if not any(os.path.exists(os.path.join(ROOT, p + m)) for p in prefixes):
            missing.append(m)
    assert not missing, missing
    for name in re.findall(r"::(test_[a-z0-9_]+)", text):      # named tests exist somewhere under tests/
        hits = [f for f in os.listdir(os.path.join(ROOT, "tests")) if f.endswith(".py")
                and f"def {name}" in open(os.path.join(ROOT, "tests", f)).read()]
        assert hits, name


def test_ncu_table_unit_conversion():
    import importlib.util
    spec = importlib.util.spec_from_file_location("ncu_table", os.path.join(ROOT, "scripts", "ncu_table.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    assert mod.num("53.6", "us", "time") == "53.6" and mod.num("1.5", "ms", "time") == "1500"
    assert mod.num("12.5", "Mbyte", "bytes") == "12.5" and mod.num("2,048", "Kbyte", "bytes") == "2.0"
    assert mod.num("n/a", "", "") == "n/a"

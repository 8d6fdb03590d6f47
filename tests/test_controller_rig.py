"""Drop-in proof without a cluster: the reference's own deployer renders
the PyTorchJob for our plugin; the rig (tests/rig/render_manifest.py) only fakes the kubernetes
client, the settings object and the logger.  Skipped where /root/reference is not mounted."""
import json
import os
import re
import subprocess
import sys

import pytest

REF = "/root/reference"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PLUGIN = os.path.join(ROOT, "finetune-controller_b200", "plugin", "b200_lora.py")
RIG = os.path.join(ROOT, "tests", "rig", "render_manifest.py")

pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "app")), reason="reference tree not mounted")


def render(n_gpu, targs=None):
    out = subprocess.run([sys.executable, RIG, REF, PLUGIN, str(n_gpu), json.dumps(targs or {})],
                         capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr[-2000:]
    line = [l for l in out.stdout.splitlines() if l.startswith("MANIFEST_JSON=")][-1]
    return json.loads(line[len("MANIFEST_JSON="):])


def test_manifest_for_an_8_gpu_job():
    sub = render(8, {"epochs": 2, "lora_r": 16, "attention": "own"})
    assert (sub["group"], sub["version"], sub["plural"]) == ("kubeflow.org", "v1", "pytorchjobs")
    body = sub["body"]
    assert body["kind"] == "PyTorchJob" and body["metadata"]["name"] == "b200-lora-0a1b2c3d"
    # DNS-1035 label: the operator names the "<job>-master-0" Service and pod after it
    assert re.fullmatch(r"[a-z]([-a-z0-9]*[a-z0-9])?", body["metadata"]["name"])
    assert body["spec"]["runPolicy"]["suspend"] is True            # Kueue local queue
    assert body["metadata"]["labels"]["kueue.x-k8s.io/queue-name"] == "finetune-queue"
    assert body["metadata"]["labels"]["job.accelerators"] == "8"
    specs = body["spec"]["pytorchReplicaSpecs"]
    assert set(specs) == {"Master"}                                 # one box: cluster_nodes = 1
    pod = specs["Master"]["template"]["spec"]
    main, sidecar = pod["containers"]
    assert main["name"] == "pytorch" and sidecar["name"] == "s3-sync"
    cmd = main["command"]
    assert cmd[:2] == ["/bin/bash", "-c"]
    assert cmd[2].startswith("python -m b200ft.train ") and cmd[2].endswith("&& touch /data/artifacts/done.txt")
    for flag in ("--gpus=8", "--epochs=2", "--lora_r=16", "--attention=own", "--nodes=1", "--dataset_path=/data/dataset",
                 "--checkpoint_path=/data/artifacts"):
        assert flag in cmd[2], flag
    res = main["resources"]
    assert res["requests"]["nvidia.com/gpu"] == 8 and res["limits"]["nvidia.com/gpu"] == 8
    assert res["requests"]["cpu"] == 16                             # the plugin's resources replace worker defaults
    assert {"name": "NCCL_DEBUG", "value": "INFO"} in main["env"]
    mounts = {m["name"]: m["mountPath"] for m in main["volumeMounts"]}
    assert mounts == {"model-checkpoint-volume": "/data/artifacts", "dataset-volume": "/data/dataset", "dshm": "/dev/shm"}
    assert pod["tolerations"][0]["key"] == "nvidia.com/gpu"
    assert pod["initContainers"] == []                              # no dataset uploaded
    # the sidecar's sync patterns pick up what the payload writes
    assert "--include '*.csv'" in sidecar["args"][1] and "--include '*.pt'" in sidecar["args"][1]
    assert "--include '*.safetensors'" in sidecar["args"][1] and "state" not in sidecar["args"][1]


def test_manifest_single_gpu_and_entrypoint_accepts_the_command():
    sub = render(1)
    main = sub["body"]["spec"]["pytorchReplicaSpecs"]["Master"]["template"]["spec"]["containers"][0]
    assert main["resources"]["limits"]["nvidia.com/gpu"] == 1
    sys.path.insert(0, os.path.join(ROOT, "finetune-controller_b200"))
    from b200ft.train import build_parser
    shell = main["command"][2]
    argv = shell.split(" && touch ")[0].split()[3:]
    ns = build_parser().parse_args(argv)
    assert ns.gpus == 1 and ns.model == "gpt2-124m" and ns.checkpoint_path == "/data/artifacts"

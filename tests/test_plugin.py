"""The drop-in boundary: our plugin file against the reference's OWN plugin machinery
(app/models/base/finetuning.py, app/models/model_loader.py, app/jobs/registered_models.py),
imported from /root/reference when it is present (it is in the build container; these are
`not gpu` tests).  Mirrors how the controller uses a plugin: discovery, registration by name,
default construction, form schema, validation/coercion on submit, run_cmd()."""
import os
import re
import shutil
import sys

import pytest

REF = "/root/reference"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PLUGIN = os.path.join(ROOT, "finetune-controller_b200", "plugin", "b200_lora.py")

pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "app")), reason="reference tree not mounted")


@pytest.fixture(scope="module")
def plugin_cls(tmp_path_factory):
    if REF not in sys.path:
        sys.path.insert(0, REF)
    from app.models.model_loader import load_models_from_directory
    d = tmp_path_factory.mktemp("custom")
    shutil.copy(PLUGIN, d / "b200_lora.py")
    models = load_models_from_directory(str(d))
    assert "B200LoRA" in models, "plugin was not discovered by the reference loader"
    return models["B200LoRA"]


def test_registers_under_its_name_like_registered_models_does(plugin_cls):
    name = plugin_cls.model_fields.get("name").get_default()      # registered_models.py:28
    assert name == "B200_LoRA" and len(name) >= 4 and re.fullmatch(r"[a-zA-Z0-9._@]+", name)   # finetuning.py:57
    k8s = name.lower().replace("_", "-") + "-abcdef12"              # main.py:422 job id
    # the operator derives "<job>-master-0" Service / pod host names: DNS-1035 label, no dots
    assert re.fullmatch(r"[a-z]([-a-z0-9]*[a-z0-9])?", k8s) and len(k8s + "-master-0") <= 63


def test_default_construction_and_form_schema(plugin_cls):
    inst = plugin_cls()                                            # main.py:1344-1347
    props = inst.training_arguments.model_json_schema()["properties"]   # main.py:265-269
    assert {"model", "lora_r", "lora_alpha", "seq_len", "batch_size", "lr", "epochs"} <= set(props)
    for k, v in props.items():
        assert "type" in v and "default" in v and "description" in v, k
    assert inst.checkpoint_mount == "/data/artifacts" and inst.dataset_mount == "/data/dataset"
    assert inst.task.value in ("regression", "classification", "multitask_classification")


def test_submit_validation_coerces_and_ignores_unknown(plugin_cls):
    inst = plugin_cls(training_arguments={"lora_r": "16", "lr": "0.001", "not_a_field": 1})   # main.py:444-448
    assert inst.training_arguments.lora_r == 16 and inst.training_arguments.lr == 1e-3
    assert not hasattr(inst.training_arguments, "not_a_field")
    import pydantic
    with pytest.raises(pydantic.ValidationError):
        plugin_cls(training_arguments={"lora_r": "sixteen"})
    with pytest.raises(pydantic.ValidationError):
        plugin_cls(accelerator_count=0)


def test_run_cmd_shape_and_controller_suffix(plugin_cls):
    inst = plugin_cls(accelerator_count=8, training_arguments={"epochs": 2})
    cmd = inst.run_cmd()
    assert cmd[:2] == ["/bin/bash", "-c"] and isinstance(cmd[-1], str)
    assert "--dataset_path=/data/dataset" in cmd[-1] and "--checkpoint_path=/data/artifacts" in cmd[-1]
    assert "--gpus=8" in cmd[-1] and "--epochs=2" in cmd[-1]
    assert inst.command[-1] == "python -m b200ft.train", "run_cmd must not mutate the class default"
    # PyTorchJobDeployer.py:30-32 appends the done marker to the last element
    cmd[-1] = f"{cmd[-1]} && touch {inst.checkpoint_mount}/done.txt"
    assert cmd[-1].endswith("done.txt")


def test_entrypoint_parses_exactly_what_run_cmd_emits(plugin_cls):
    sys.path.insert(0, os.path.join(ROOT, "finetune-controller_b200"))
    from b200ft.train import build_parser
    inst = plugin_cls(accelerator_count=2, training_arguments={"lora_r": 16, "save_model": True, "synthetic": False})
    argv = inst.run_cmd()[-1].split()[3:]       # drop "python -m b200ft.train"
    ns = build_parser().parse_args(argv)         # argparse raises SystemExit on an unknown flag
    assert ns.lora_r == 16 and ns.gpus == 2 and ns.save_model and not ns.synthetic
    assert ns.dataset_path == "/data/dataset" and ns.checkpoint_path == "/data/artifacts"
    ta = inst.training_arguments
    for k in type(ta).model_fields:
        assert getattr(ns, k) == getattr(ta, k), k


def test_every_declared_argument_reaches_the_entrypoint(plugin_cls):
    """A non-default value for EVERY B200LoRAConfig field must come out of run_cmd() and be parsed back
    to the same value by the entrypoint's argparse: no declared argument can be dropped on the way."""
    sys.path.insert(0, os.path.join(ROOT, "finetune-controller_b200"))
    from b200ft.train import build_parser
    changed = {"model": "llama-3.2-1b", "mode": "full", "attention": "own", "lora_r": 4, "lora_alpha": 3.5,
               "seq_len": 256, "batch_size": 3, "epochs": 7, "steps_per_epoch": 11, "lr": 3e-5, "weight_decay": 0.2,
               "beta1": 0.8, "beta2": 0.95, "eps": 1e-6, "max_grad_norm": 1.5, "seed": 42, "log_interval": 3,
               "save_interval": -1, "save_model": False, "base_weights": "hf/gpt2", "synthetic": False}
    fields = plugin_cls.model_fields["training_arguments"].annotation.model_fields
    assert set(changed) == set(fields), "extend this test when a field is added to B200LoRAConfig"
    inst = plugin_cls(accelerator_count=4, cluster_nodes=2, training_arguments=changed)
    for k, v in changed.items():
        assert fields[k].get_default() != v, f"{k}: pick a non-default value"
    shell = inst.run_cmd()[-1]
    assert "--attention=own" in shell and "--nodes=2" in shell and "--gpus=4" in shell
    ns = build_parser().parse_args(shell.split()[3:])
    for k, v in changed.items():
        assert getattr(ns, k) == v, k
    assert ns.nodes == 2 and ns.gpus == 4


def test_multi_pod_world_is_pods_times_gpus():
    """cluster_nodes > 1: the operator's env counts pods (reference PyTorchJobDeployer.py:186-249); the
    entrypoint's ranks are pod x gpus + local."""
    sys.path.insert(0, os.path.join(ROOT, "finetune-controller_b200"))
    from b200ft.train import build_parser, pod_world
    ns = build_parser().parse_args(["--gpus=8", "--nodes=2"])
    env = {"RANK": "1", "WORLD_SIZE": "2", "MASTER_ADDR": "job-master-0", "MASTER_PORT": "23456"}
    assert pod_world(ns, env) == (8, 16, "job-master-0", 23456)
    assert pod_world(build_parser().parse_args(["--gpus=4"]), {"RANK": "3", "WORLD_SIZE": "9"}) == (0, 4, "127.0.0.1", None)
    with pytest.raises(SystemExit):
        pod_world(ns, {"RANK": "0"})                                   # operator env missing
    with pytest.raises(SystemExit):
        pod_world(ns, dict(env, WORLD_SIZE="3"))                       # disagrees with cluster_nodes


def test_type_lock_rejects_a_retyped_field():
    if REF not in sys.path:
        sys.path.insert(0, REF)
    from app.models.base.finetuning import BaseFineTuneModel
    with pytest.raises(TypeError):
        class Bad(BaseFineTuneModel):                                # finetuning.py:110-145
            command: str = "x"

            def run_cmd(self):
                return []


def test_resources_and_patterns_fit_the_deployer(plugin_cls):
    inst = plugin_cls()
    # PyTorchJobDeployer.py:41-51 iterates the resources model as (field, value) pairs
    d = dict(inst.resources)
    assert set(d) == {"requests", "limits"} and "cpu" in d["requests"]
    import fnmatch
    for produced in ("metrics.csv", "adapter.pt", "adapter_config.json", "adapter_model.safetensors", "model.pt"):
        assert any(fnmatch.fnmatch(produced, pat) for pat in inst.store_asset_patterns), produced
    # the resume point (up to 96 GB of optimizer state in full-parameter mode) must NOT be synced to S3 every
    # AWS_JOB_SYNC_INTERVAL by the sidecar (PyTorchJobDeployer.py:122-139)
    sys.path.insert(0, os.path.join(ROOT, "finetune-controller_b200"))
    from b200ft.train import STATE_FILE
    for kept_local in (STATE_FILE, STATE_FILE + ".tmp", STATE_FILE + ".P.000000100", STATE_FILE + ".vn.000000100"):
        assert not any(fnmatch.fnmatch(kept_local, pat) for pat in inst.store_asset_patterns), kept_local

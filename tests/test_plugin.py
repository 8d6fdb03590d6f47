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
    assert name == "B200.LoRA" and len(name) >= 4 and re.fullmatch(r"[a-zA-Z0-9._@]+", name)
    k8s = name.lower().replace("_", "-") + "-abcdef12"              # main.py:422 job id
    assert re.fullmatch(r"[a-z0-9]([-a-z0-9.]*[a-z0-9])?", k8s)


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
    for produced in ("metrics.csv", "adapter.pt", "adapter_config.json"):
        assert any(fnmatch.fnmatch(produced, pat) for pat in inst.store_asset_patterns), produced

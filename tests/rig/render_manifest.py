"""Controller-side rig: runs the REFERENCE's own PyTorchJobDeployer.create_pytorch_job
(reference app/jobs/kubeflow/PyTorchJobDeployer.py:20-262) on our plugin with the cluster
dependencies replaced by in-process fakes, and prints the PyTorchJob manifest it would have
submitted as JSON.  Nothing of the reference is modified or copied: the fakes stand in for the
`kubernetes` client library, the .env-backed settings object and the colour logger, none of
which exist in this container.

usage: python render_manifest.py <reference_root> <plugin_file> <accelerator_count> [json training args]
"""
import json
import os
import shutil
import sys
import tempfile
import types


def main():
    ref_root, plugin_file, n_gpu = sys.argv[1], sys.argv[2], int(sys.argv[3])
    targs = json.loads(sys.argv[4]) if len(sys.argv) > 4 else {}
    work = tempfile.mkdtemp()
    cfg_path = os.path.join(work, "config.json")
    with open(cfg_path, "w") as f:
        f.write("""{
  // the worker entry INTEGRATION.md proposes
  "default_queue": "finetune-queue",
  "workers": [
    {"name": "cpu", "local_queue": "finetune-queue", "defaults": {"resources": {"requests": {"cpu": 2, "memory": "2Gi"}}}},
    {"name": "b200-8x", "local_queue": "finetune-queue",
     "defaults": {"accelerators": {"nvidia.com/gpu": 8}, "resources": {"requests": {"cpu": 16, "memory": "64Gi"}}},
     "tolerations": [{"key": "nvidia.com/gpu", "value": "b200", "effect": "NoSchedule"}]}
  ]
}""")
    submitted = {}

    class FakeCustomObjectsApi:
        def create_namespaced_custom_object(self, group, version, namespace, plural, body):
            submitted.update(group=group, version=version, namespace=namespace, plural=plural, body=body)
            return {"metadata": {"name": body["metadata"]["name"]}}

    k8s = types.ModuleType("kubernetes")
    k8s.client = types.ModuleType("kubernetes.client")
    k8s.client.CustomObjectsApi = FakeCustomObjectsApi
    k8s.client.CoreV1Api = type("CoreV1Api", (), {})
    sys.modules["kubernetes"], sys.modules["kubernetes.client"] = k8s, k8s.client

    fake_cfg = types.ModuleType("app.core.config")
    fake_cfg.settings = types.SimpleNamespace(
        CONFIGURATION_FILE=cfg_path, MONGODB_DATABASE="rig", AWS_SECRET_NAME="aws-secret", AWS_REGION="us-east-1",
        S3_BUCKET_NAME="rig-bucket", AWS_JOB_SYNC_INTERVAL=60, NAMESPACE="finetune")
    fake_log = types.ModuleType("app.utils.logging_config")
    fake_log.setup_logging = lambda *a, **k: None

    sys.path.insert(0, ref_root)
    import app.core  # noqa: F401  (real package)
    import app.utils  # noqa: F401
    sys.modules["app.core.config"] = fake_cfg
    sys.modules["app.utils.logging_config"] = fake_log

    from app.models.model_loader import load_models_from_directory
    custom = os.path.join(work, "custom")
    os.makedirs(custom)
    shutil.copy(plugin_file, custom)
    cls = load_models_from_directory(custom)["B200LoRA"]

    from app.jobs.kubeflow.PyTorchJobDeployer import PyTorchJobDeployer
    from app.schemas.jobs_schemas import JobInput

    model = cls(training_arguments=targs, accelerator_count=n_gpu)            # main.py:444-448
    name = cls.model_fields["name"].get_default()
    job = JobInput(user_id="rig-user", job_name="rig", model_name=name, model=model, device="b200-8x",
                   arguments=model.training_arguments.model_dump(), s3_uri="",
                   s3_artifacts_uri="s3://rig-bucket/finetune_jobs/rig-user/job/artifacts",
                   job_id=name.lower().replace("_", "-") + "-0a1b2c3d")         # main.py:422
    PyTorchJobDeployer(namespace="finetune").create_pytorch_job(job)
    print("MANIFEST_JSON=" + json.dumps(submitted))


if __name__ == "__main__":
    main()

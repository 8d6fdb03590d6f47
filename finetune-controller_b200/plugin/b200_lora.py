"""Model plugin for acceleratedscience/finetune-controller: drop this ONE file into the
controller's ``app/models/custom/`` directory.

It is discovered by ``load_models_from_directory`` (reference app/models/model_loader.py:14-45),
registered under its ``name`` default (reference app/jobs/registered_models.py:20-37), shown
in the UI through ``training_arguments.model_json_schema()`` (reference app/main.py:265-269),
validated on submit (reference app/main.py:439-471) and turned into the ``pytorch``
container's command by ``run_cmd()`` (reference app/jobs/kubeflow/PyTorchJobDeployer.py:28-32).
Structure follows the shipped exemplar (reference app/models/examples/mnist.py:13-99).

This module runs inside the API server, so it imports pydantic-level code only — no torch.
The command it emits starts ``python -m b200ft.train`` (finetune-controller_b200/b200ft/train.py),
which runs the B200 kernels.
"""
from app.models.base.finetuning import (
    BaseFineTuneModel,
    Field,
    TrainingArguments,
    TrainingDataset,
    TrainingFramework,
    TrainingResources,
    TrainingTask,
)


class B200LoRAConfig(TrainingArguments):
    """Hyper-parameters of the B200-native LoRA fine-tune job (one type + default each)."""

    model: str = Field(default="gpt2-124m", description="Architecture: gpt2-124m | llama-3.2-1b | llama-3-8b (gpt2-tiny, llama-tiny for smoke runs)")
    mode: str = Field(default="lora", description="lora: adapters on c_attn (GPT-2) or q_proj + v_proj (Llama); full: every tensor trainable (Llama family; fp32 master + Adam moments, 8B fits one 180 GB B200)")
    attention: str = Field(default="auto", description="own: this library's tcgen05 attention kernels (gpt2-124m, llama-3.2-1b, llama-3-8b); sdpa: cuDNN flash attention through torch; auto: own wherever it is built")
    lora_r: int = Field(default=8, description="LoRA rank (<= 16)")
    lora_alpha: float = Field(default=16.0, description="LoRA alpha; the branch is scaled by alpha / r")
    seq_len: int = Field(default=1024, description="Tokens per sequence")
    batch_size: int = Field(default=32, description="Sequences per GPU per optimizer step")
    epochs: int = Field(default=1, description="Number of epochs")
    steps_per_epoch: int = Field(default=100, description="Optimizer steps per epoch (synthetic data) or cap")
    lr: float = Field(default=1e-4, description="AdamW learning rate")
    weight_decay: float = Field(default=0.01, description="AdamW decoupled weight decay")
    beta1: float = Field(default=0.9, description="AdamW beta1")
    beta2: float = Field(default=0.999, description="AdamW beta2")
    eps: float = Field(default=1e-8, description="AdamW epsilon")
    max_grad_norm: float = Field(default=0.0, description="Global gradient-norm clip; 0 disables")
    seed: int = Field(default=0, description="Seed for weight init and synthetic data")
    log_interval: int = Field(default=10, description="Steps between metric rows / log lines")
    save_interval: int = Field(default=0, description="Steps between resume points (resume.state, not uploaded); 0 = once per epoch, -1 = never")
    save_model: bool = Field(default=True, description="Write adapter_model.safetensors (PEFT key names) + adapter.pt + adapter_config.json (lora) or model.pt + config.json (full) at the end")
    base_weights: str = Field(default="", description="Pretrained base weights: a Hugging Face *.safetensors file or directory (or a .pt state dict), relative to the dataset mount; empty = random init")
    synthetic: bool = Field(default=True, description="Use synthetic token ids instead of the mounted dataset")


class B200LoRA(BaseFineTuneModel):
    """Job spec: GPT-2-class LoRA fine-tuning on B200 GPUs with hand-written sm_100a kernels."""

    # the controller derives the PyTorchJob name as name.lower().replace("_", "-") + "-" + uuid8
    # (reference app/main.py:422) and the operator derives "<job>-master-0" Service / pod host names
    # from it: "b200-lora-xxxxxxxx" is a DNS-1035 label, a dotted name would not be
    name: str = "B200_LoRA"
    inference_name: str | None = None
    description: str = "GPT-2 LoRA fine-tune on NVIDIA B200 (tcgen05 GEMMs, fused CE, fused AdamW)"
    project_url: str = ""
    image: str = "registry.local/b200ft:latest"
    command: list[str] = [
        "/bin/bash",
        "-c",
        "python -m b200ft.train",
    ]
    framework: TrainingFramework = TrainingFramework.PYTORCH
    # the reference enum has no causal-LM member (finetuning.py:8-11); next-token prediction is
    # a classification over the vocabulary
    task: TrainingTask = TrainingTask.CLASSIFICATION
    dataset_info: TrainingDataset = TrainingDataset(
        description="Optional tokens.pt / tokens.bin (int token ids) under the dataset mount; synthetic otherwise",
        dataset_required=False,
    )
    device_types: list[str] = ["b200-8x"]
    resources: TrainingResources = TrainingResources(
        requests={"cpu": 16, "memory": "64Gi"}, limits={"cpu": 32, "memory": "128Gi"}
    )
    accelerator_count: int = Field(
        default=1,
        ge=1,
        description="Number of gpu devices to use for training per worker",
    )
    # the base patterns (reference app/models/base/finetuning.py:94-97) plus the PEFT adapter file
    store_asset_patterns: list[str] = Field(
        default=["*.json", "*.yaml", "*.csv", "*.pt", "*.ckpt", "*.safetensors"],
        description="Pattern match a list of files to store.",
    )
    promotion_path: str = Field(
        default="language/gpt2/b200_lora",
        description="s3 path to upload artifacts. Based on Inference path `domain/algorithm_name/algorithm_application`",
    )

    training_arguments: B200LoRAConfig = B200LoRAConfig()

    def run_cmd(self) -> list[str]:
        cmd = self.command.copy()
        ta = self.training_arguments
        args = []
        # every declared argument goes on the command line (a field added to B200LoRAConfig can
        # not be forgotten here): booleans as bare flags like the exemplar (reference
        # app/models/examples/mnist.py:75-99), the rest as --key=value; empty strings are left out
        for key in type(ta).model_fields:
            val = getattr(ta, key)
            if isinstance(val, bool):
                if val:
                    args.append(f"--{key}")
            elif val != "":
                args.append(f"--{key}={val}")
        args.append(f"--gpus={self.accelerator_count}")
        # Master + Workers of the PyTorchJob (reference PyTorchJobDeployer.py:186-249): the entrypoint
        # builds its world as pods x gpus from the operator's per-pod RANK / WORLD_SIZE
        args.append(f"--nodes={self.cluster_nodes}")
        # the two mounts every payload must honour (reference mnist.py:93-95)
        args.append(f"--dataset_path={self.dataset_mount}")
        args.append(f"--checkpoint_path={self.checkpoint_mount}")
        cmd[-1] += " " + " ".join(args)
        return cmd


if __name__ == "__main__":
    from pprint import pprint

    job = B200LoRA(training_arguments={"epochs": 2, "lora_r": "16"})
    pprint(job.model_dump())
    print(job.run_cmd())

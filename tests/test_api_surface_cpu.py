"""Every keyword argument the reference's scripts pass to the HF / PEFT / quantisation entry points this framework mirrors is
accepted by the mirror (an AST scan of /root/reference; skipped where that tree is absent)."""
import ast
import dataclasses
import glob
import inspect
import os

import pytest

REF = "/root/reference"


def _accepted(obj):
    if dataclasses.is_dataclass(obj):
        return {f.name for f in dataclasses.fields(obj)}, False
    params = inspect.signature(obj).parameters
    return set(params), any(p.kind == p.VAR_KEYWORD for p in params.values())


def test_reference_call_sites_are_accepted():
    if not os.path.isdir(REF):
        pytest.skip("reference tree not present")
    from llm_in_practise_b200 import peft, quant
    from llm_in_practise_b200.data import datasets as D
    from llm_in_practise_b200.inference.generate import GenerationConfig
    from llm_in_practise_b200.io import hf
    from llm_in_practise_b200.trainer.args import TrainingArguments
    from llm_in_practise_b200.trainer.trainer import Trainer
    mirrors = {"TrainingArguments": TrainingArguments, "LoraConfig": peft.LoraConfig, "BitsAndBytesConfig": hf.BitsAndBytesConfig,
               "Trainer": Trainer, "DataCollatorForLanguageModeling": D.DataCollatorForLanguageModeling,
               "DataCollatorWithPadding": D.DataCollatorWithPadding, "get_peft_model": peft.get_peft_model,
               "QuantizeConfig": quant.QuantizeConfig, "GPTQModifier": quant.GPTQModifier, "AWQModifier": quant.AWQModifier,
               "QuantizationModifier": quant.QuantizationModifier, "oneshot": quant.oneshot, "GenerationConfig": GenerationConfig}
    files = [f for pat in ("Fine-Tuning/*.py", "HF_Basics/*.py", "Scripts/**/*.py", "Quantization/**/*.py")
             for f in glob.glob(os.path.join(REF, pat), recursive=True)]
    assert len(files) > 20
    calls, refused = 0, []
    for f in files:
        try:
            tree = ast.parse(open(f, encoding="utf-8").read())
        except SyntaxError:
            continue
        for node in ast.walk(tree):
            if not isinstance(node, ast.Call):
                continue
            name = getattr(node.func, "id", getattr(node.func, "attr", None))
            if name not in mirrors:
                continue
            calls += 1
            names, varkw = _accepted(mirrors[name])
            bad = [k.arg for k in node.keywords if k.arg and k.arg not in names]
            if bad and not varkw:
                refused.append((os.path.relpath(f, REF), name, bad))
    assert calls > 30, calls
    assert not refused, refused


def test_distributed_training_call_sites_are_accepted():
    """Same scan over LLM_Distributed_Trainning/**: DDP / FSDP / fully_shard / MixedPrecision / DistributedSampler /
    deepspeed.initialize keyword arguments."""
    if not os.path.isdir(REF):
        pytest.skip("reference tree not present")
    from llm_in_practise_b200 import compat as C
    mirrors = {"DDP": C.DDP, "DistributedDataParallel": C.DDP, "FSDP": C.FSDP, "FullyShardedDataParallel": C.FSDP,
               "fully_shard": C.fully_shard, "MixedPrecision": C.MixedPrecision, "MixedPrecisionPolicy": C.MixedPrecisionPolicy,
               "CPUOffload": C.CPUOffload, "CPUOffloadPolicy": C.CPUOffloadPolicy, "DistributedSampler": C.DistributedSampler,
               "DeepSpeedCPUAdam": C.DeepSpeedCPUAdam, "init_device_mesh": C.init_device_mesh, "StateDictOptions": C.StateDictOptions,
               "transformer_auto_wrap_policy": C.transformer_auto_wrap_policy, "get_state_dict": C.get_state_dict}
    from llm_in_practise_b200.parallel.zero import initialize
    mirrors["initialize"] = initialize
    calls, refused = 0, []
    for f in glob.glob(os.path.join(REF, "LLM_Distributed_Trainning/**/*.py"), recursive=True):
        try:
            tree = ast.parse(open(f, encoding="utf-8").read())
        except SyntaxError:
            continue
        for node in ast.walk(tree):
            if isinstance(node, ast.Call):
                name = getattr(node.func, "id", getattr(node.func, "attr", None))
                if name in mirrors:
                    calls += 1
                    names, varkw = _accepted(mirrors[name])
                    bad = [k.arg for k in node.keywords if k.arg and k.arg not in names]
                    if bad and not varkw:
                        refused.append((os.path.relpath(f, REF), name, bad))
    assert calls > 30, calls
    assert not refused, refused


def test_methods_the_scripts_call_exist():
    """``model.<x>(...)`` / ``trainer.<x>(...)`` / ``peft_model.<x>(...)`` in the reference's scripts: each name exists on the
    object that plays that role here (``model`` is a causal LM, a PEFT wrapper or a GPTQModel depending on the script)."""
    if not os.path.isdir(REF):
        pytest.skip("reference tree not present")
    from llm_in_practise_b200.models.classifier import EncoderClassifier
    from llm_in_practise_b200.models.qwen3 import Qwen3ForCausalLM
    from llm_in_practise_b200.peft import PeftModel
    from llm_in_practise_b200.quant import GPTQModel
    from llm_in_practise_b200.trainer.trainer import Trainer
    roles = {"model": (Qwen3ForCausalLM, PeftModel, GPTQModel, EncoderClassifier), "peft_model": (PeftModel, Qwen3ForCausalLM),
             "trainer": (Trainer,)}
    files = [f for pat in ("Fine-Tuning/*.py", "HF_Basics/*.py", "Scripts/**/*.py", "Quantization/**/*.py")
             for f in glob.glob(os.path.join(REF, pat), recursive=True)]
    used = {k: set() for k in roles}
    for f in files:
        try:
            tree = ast.parse(open(f, encoding="utf-8").read())
        except SyntaxError:
            continue
        for node in ast.walk(tree):
            if isinstance(node, ast.Call) and isinstance(node.func, ast.Attribute) and isinstance(node.func.value, ast.Name) \
                    and node.func.value.id in roles:
                used[node.func.value.id].add(node.func.attr)
    assert {"generate", "save_pretrained", "print_trainable_parameters"} <= used["model"] and "train" in used["trainer"]
    missing = [(role, name) for role, names in used.items() for name in sorted(names)
               if not any(hasattr(cls, name) for cls in roles[role])]
    assert not missing, missing

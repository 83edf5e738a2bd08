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

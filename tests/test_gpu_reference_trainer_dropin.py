"""The drop-in claim of INTEGRATION.md §1, proven by the reference's OWN loop: the unmodified ``trainers.Trainer``
(``trainers.py:9-158``) and the unmodified ``main.py --do_eval`` (``main.py:37-48``) run with
``MODEL_DICT["bsarec"] = bsarec_b200.BSARecModel`` (``model/__init__.py:10-11``, ``main.py:31``) and nothing else changed.

Needs the reference sources next to a GPU: ``/root/reference/src`` (build container — no GPU there) or
``baseline/_ref/src`` staged by ``baseline/stage_ref.sh`` (git-ignored, travels with gpurun).  Skipped when neither is
present; the committed evidence run is ``profiles/r2a_dropin_reference_trainer.log``."""
import argparse
import logging
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from conftest import ROOT

pytestmark = pytest.mark.gpu

CANDIDATES = [os.environ.get("BSAREC_REF_SRC", ""), "/root/reference/src", os.path.join(ROOT, "baseline", "_ref", "src")]
REF_SRC = next((p for p in CANDIDATES if p and os.path.isfile(os.path.join(p, "trainers.py"))), None)
needs_ref = pytest.mark.skipif(REF_SRC is None, reason="reference sources not staged (baseline/stage_ref.sh)")
LOG_258 = ["0.0736", "0.0523", "0.1008", "0.0611", "0.1373", "0.0703"]       # BSARec_Beauty_best.log:258


def _ref_args(tmp_path, **kw):
    a = argparse.Namespace(
        data_dir=os.path.join(REF_SRC, "data") + "/", output_dir=str(tmp_path) + "/", data_name="Beauty", do_eval=True,
        load_model="BSARec_Beauty_best", train_name="dropin", num_items=10, num_users=10, lr=0.001, batch_size=256, epochs=1,
        no_cuda=False, log_freq=1, patience=10, num_workers=0, seed=42, weight_decay=0.0, adam_beta1=0.9, adam_beta2=0.999,
        gpu_id="0", variance=5, model_type="BSARec", max_seq_length=50, hidden_size=64, num_hidden_layers=2, hidden_act="gelu",
        num_attention_heads=1, attention_probs_dropout_prob=0.5, hidden_dropout_prob=0.5, initializer_range=0.02, c=5,
        alpha=0.7)
    for k, v in kw.items():
        setattr(a, k, v)
    return a


@needs_ref
def test_reference_trainer_runs_on_the_swapped_class(tmp_path, cuda_device):
    """trainers.Trainer as it lies: test() through predict()/predict_full/numpy top-20 reproduces log:258, then its train
    branch (calculate_loss -> zero_grad -> backward -> Adam.step -> loss.item()) runs on our class and lowers the loss."""
    sys.path.insert(0, REF_SRC)
    try:
        import model as ref_models                      # the reference's package
        from trainers import Trainer as RefTrainer
        import dataset as ref_dataset
        import bsarec_b200
        ref_models.MODEL_DICT["bsarec"] = bsarec_b200.BSARecModel          # <- the one change of INTEGRATION.md §1
        args = _ref_args(tmp_path)
        seq_dic, max_item, num_users = ref_dataset.get_seq_dic(args)
        args.item_size, args.num_users = max_item + 1, num_users + 1
        args.valid_rating_matrix, args.test_rating_matrix = ref_dataset.get_rating_matrix(args.data_name, seq_dic, max_item)
        train_dl, eval_dl, test_dl = ref_dataset.get_dataloder(args, seq_dic)
        model = ref_models.MODEL_DICT[args.model_type.lower()](args=args)
        assert type(model) is bsarec_b200.BSARecModel
        trainer = RefTrainer(model, train_dl, eval_dl, test_dl, args, logging.getLogger("dropin"))
        trainer.load(os.path.join(REF_SRC, "output", "BSARec_Beauty_best.pt"))
        scores, info = trainer.test(0)
        assert ["%.4f" % v for v in scores] == LOG_258, info
        # train branch on a few of the reference's own batches
        batches = []
        for i, b in enumerate(train_dl):
            batches.append(b)
            if i == 5:
                break
        losses = []
        orig_item = torch.Tensor.item
        trainer.model.train()
        for b in batches * 2:
            user_ids, input_ids, answers, neg_answer, same_target = (t.to(trainer.device) for t in b)
            loss = trainer.model.calculate_loss(input_ids, answers, neg_answer, same_target, user_ids)
            trainer.optim.zero_grad()
            loss.backward()
            trainer.optim.step()
            losses.append(loss.item())
        assert all(np.isfinite(losses))
        trainer.iteration(0, batches, train=True)          # the reference's own loop body, unmodified
        assert orig_item is torch.Tensor.item
    finally:
        sys.path.remove(REF_SRC)
        for m in ("model", "trainers", "dataset", "metrics", "utils"):
            sys.modules.pop(m, None)
        for m in [k for k in sys.modules if k.startswith("model.")]:
            sys.modules.pop(m, None)


@needs_ref
def test_reference_main_do_eval_with_the_swapped_class(tmp_path, cuda_device):
    """`python main.py --do_eval` of the reference, run as a script in a subprocess; the only difference is a sitecustomize
    that performs the MODEL_DICT swap before main.py is imported."""
    import shutil
    shutil.copy(os.path.join(REF_SRC, "output", "BSARec_Beauty_best.pt"), os.path.join(tmp_path, "BSARec_Beauty_best.pt"))
    hook = os.path.join(tmp_path, "hook")
    os.makedirs(hook)
    with open(os.path.join(hook, "sitecustomize.py"), "w") as f:
        f.write(f"import sys\nsys.path.insert(0, {ROOT!r})\nsys.path.insert(0, {REF_SRC!r})\n"
                "import model, bsarec_b200\nmodel.MODEL_DICT['bsarec'] = bsarec_b200.BSARecModel\n")
    env = dict(os.environ, PYTHONPATH=hook + os.pathsep + os.environ.get("PYTHONPATH", ""))
    cmd = [sys.executable, os.path.join(REF_SRC, "main.py"), "--data_dir", os.path.join(REF_SRC, "data") + "/",
           "--output_dir", str(tmp_path) + "/", "--data_name", "Beauty", "--alpha", "0.7", "--c", "5",
           "--num_attention_heads", "1", "--load_model", "BSARec_Beauty_best", "--do_eval", "--num_workers", "0"]
    r = subprocess.run(cmd, capture_output=True, text=True, cwd=REF_SRC, env=env, timeout=900)
    out = r.stdout + r.stderr
    assert r.returncode == 0, out[-3000:]
    for s in LOG_258:
        assert s in out, out[-3000:]
    assert "TFLayerNorm" in out          # main.py:32 logs the model: the tree printed is bsarec_b200's container classes

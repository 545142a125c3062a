"""Host-side logic of the drop-in mirror that needs no GPU: precision naming, plug-in registry, error behaviour of the
scoring dispatch, module trees under the reference's parameter names."""
import numpy as np
import pytest
import torch


def test_precision_names():
    from speakerverification_b200.engine import normalize_precision
    assert normalize_precision("fp32") == "fp32"
    for alias in ("f16", "fp16", "half", "bf16"):          # 'bf16' = BASELINE.json's name for the 16-bit config
        assert normalize_precision(alias) == "f16"
    with pytest.raises(ValueError):
        normalize_precision("int8")


def test_scoring_dispatch_errors_without_touching_the_gpu():
    from speakerverification_b200 import utils as U
    emb = torch.zeros((4, 1, 8))
    idx = torch.zeros(2, dtype=torch.int32)
    with pytest.raises(ValueError, match="scoring_mode"):
        U.score_trials(emb, idx, idx, "bogus")
    with pytest.raises(ValueError, match="cohort"):
        U.score_trials(emb, idx, idx, "norm", cohorts=None)
    assert U.similarity_measure("no_such_method", emb[0], emb[1]) is None     # src/utils.py:126-132 falls through


def test_plugin_registry_and_refusals():
    from speakerverification_b200 import model as M
    assert set(M._ENCODERS) == {"ECAPA_TDNN", "RawNet2_custom", "Raw3_ECAPA"}
    base = dict(model={"name": "ECAPA_TDNN", "nOut": 192}, criterion={"name": "AAmSoftmaxAP"}, n_mels=80,
                dataloader_options={"nPerSpeaker": 2})
    with pytest.raises(RuntimeError, match="CUDA"):
        M.SpeakerEncoder(**dict(base, device="cpu"))
    from speakerverification_b200 import raw3_ecapa
    with pytest.raises(NotImplementedError, match="asteroid_filterbanks"):
        raw3_ecapa.MainModel(nOut=512, n_mels=80)
    with pytest.raises(ValueError):
        raw3_ecapa.MainModel(nOut=192, n_mels=80, svb_second_encoder=torch.nn.Identity())


def test_module_trees_carry_the_reference_names():
    """Parameter holders only (no libsvb call): key sets of the reference's checkpoints."""
    from speakerverification_b200 import ecapa_tdnn, rawnet2_custom
    from speakerverification_b200 import synth
    m = ecapa_tdnn.MainModel(nOut=192, n_mels=80, channels=[512, 512, 512, 512, 1536], input_norm=True)
    keys = set(m.state_dict())
    assert {"instance_norm.weight", "instance_norm.bias", "blocks.1.tdnn1.conv.conv.weight",
            "blocks.3.res2net_block.blocks.6.norm.norm.running_var", "asp.conv.conv.bias", "fc.conv.weight"} <= keys
    assert len(keys) == 233                                   # 231 + the two InstanceNorm parameters
    ecapa_tdnn.MainModel(nOut=192, n_mels=80, channels=[512, 512, 512, 512, 1536], activation=torch.nn.GELU)   # Raw3_ECAPA passes it
    with pytest.raises(NotImplementedError, match="GELU"):
        ecapa_tdnn.MainModel(nOut=192, n_mels=80, channels=[512, 512, 512, 512, 1536], activation=torch.nn.ReLU)
    spec = {"sample_rate": 16000, "sentence_len": 4.0}
    for agg in ("gru", "sap", "asp"):
        r = rawnet2_custom.MainModel(nOut=512, audio_spec=spec, features="raw", aggregate=agg)
        want = {k[len("__S__."):] for k in synth.rawnet2_state_dict(aggregate=agg)}
        assert set(r.state_dict()) == want, agg
    with pytest.raises(NotImplementedError):
        rawnet2_custom.MainModel(nOut=512, audio_spec=spec, features="raw", aggregate="mean")

"""Import the UNMODIFIED reference (hiimmuc/SpeakerVerification) on CPU.

Only usable where /root/reference exists (the build container); used by
``oracle/make_golden.py`` to generate the committed fixtures and by CPU tests
that are skipped when the tree is absent.  Non-arithmetic third-party modules
that are not installed here are replaced by empty stubs;
every arithmetic dependency (torch, torchaudio, numpy, sklearn) is the real one.
"""
import importlib.machinery
import os
import sys
import types

REF_ROOT = os.environ.get("SVB_REFERENCE_ROOT", "/root/reference")
WAVS = {}  # in-memory "files" served to the reference's soundfile.read


def available():
    return os.path.isdir(os.path.join(REF_ROOT, "src"))


def _stub(name):
    mod = types.ModuleType(name)
    mod.__spec__ = importlib.machinery.ModuleSpec(name, None)
    sys.modules[name] = mod
    return mod


def load():
    """Returns the reference's ``model`` and ``utils`` modules."""
    if not available():
        raise RuntimeError("reference tree not present at %s" % REF_ROOT)
    sys.dont_write_bytecode = True
    for name in ["torchsummary", "hyperpyyaml", "seaborn", "matplotlib", "matplotlib.pyplot",
                 "onnx", "onnxruntime", "soundfile", "pydub", "nnAudio", "nnAudio.features",
                 "librosa"]:
        if name not in sys.modules:
            _stub(name)
    sys.modules["torchsummary"].summary = lambda *a, **k: None
    sys.modules["hyperpyyaml"].load_hyperpyyaml = lambda *a, **k: None
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    sys.modules["pydub"].AudioSegment = object
    sys.modules["nnAudio"].features = sys.modules["nnAudio.features"]
    sys.modules["soundfile"].read = lambda path, *a, **k: (WAVS[str(path)], 16000)
    src = os.path.join(REF_ROOT, "src")
    if src not in sys.path:
        sys.path.insert(0, src)
    import model as ref_model  # noqa: E402
    import utils as ref_utils  # noqa: E402
    return ref_model, ref_utils


def reference_args(channels, n_out, sentence_len=3.0, features="melspectrogram",
                   model_name="ECAPA_TDNN", **extra):
    """Config dict the reference constructors need; `extra` keys (e.g. input_norm=True) reach
    the encoder's MainModel(**kwargs) like any other YAML key."""
    return dict(extra, 
        device="cpu", gpu=0, seed=0, save_folder="/tmp/svb_ref",
        model={"name": model_name, "nOut": n_out}, channels=list(channels),
        criterion={"name": "AAmSoftmaxAP", "margin": 0.2, "scale": 30}, nClasses=10,
        classifier={"input_size": n_out, "out_neurons": 10},
        optimizer={"name": "adam", "weight_decay": 2e-5, "lr_decay": 0.95}, lr=1e-3,
        callbacks={"name": "steplr"}, step_size=10,
        features=features, lib="torchaudio", n_mels=80, n_mfcc=80,
        sr=16000, n_fft=512, win_length=400, hop_length=160,
        audio_spec={"sample_rate": 16000, "channels": 1, "sentence_len": sentence_len,
                    "win_len": 0.025, "hop_len": 0.01},
        dataloader_options={"nPerSpeaker": 2, "num_workers": 0, "batch_size": 4, "shuffle": False},
        augment=False, augment_options={"augment_chain": []}, mixedprec=False)


def build(channels, n_out, **kw):
    """(net, handle): WrappedModel(SpeakerEncoder) and ModelHandling of the reference."""
    ref_model, _ = load()
    args = reference_args(channels, n_out, **kw)
    net = ref_model.WrappedModel(ref_model.SpeakerEncoder(**args))
    handle = ref_model.ModelHandling(net, **dict(args, T_max=10))
    net.eval()
    return net, handle, args

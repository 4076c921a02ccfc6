"""ImageNet-pretrained initialisation path (reference film_efficientnet_encoder.py:376-459): torchvision's EfficientNet-B3
state dict is zipped onto this module's tensors BY ORDER, then copied into the FiLM variant by name; ILSVRCPredictor names the
top-k classes.  No pretrained blob exists offline, so a randomly initialised torchvision network stands in for the weights —
what is being tested is the mapping, and that the two architectures compute the same function."""
import pytest
import torch

from rt1b200.models.film_efficientnet import EfficientNetB3, ILSVRCPredictor

torchvision = pytest.importorskip("torchvision")


def _randomise_bn(model, seed):
    g = torch.Generator().manual_seed(seed)
    for m in model.modules():
        if isinstance(m, torch.nn.BatchNorm2d):
            with torch.no_grad():
                m.running_mean.copy_(torch.randn(m.running_mean.shape, generator=g) * 0.1)
                m.running_var.copy_(torch.rand(m.running_var.shape, generator=g) + 0.5)
                m.weight.copy_(torch.rand(m.weight.shape, generator=g) + 0.5)
                m.bias.copy_(torch.randn(m.bias.shape, generator=g) * 0.1)


def test_torchvision_state_dict_loads_by_position_and_matches_torchvision(tmp_path):
    torch.manual_seed(0)
    tv = torchvision.models.efficientnet_b3(weights=None).eval()
    _randomise_bn(tv, 1)
    path = str(tmp_path / "efficientnetb3.pth")
    torch.save(tv.state_dict(), path)
    x = torch.rand(1, 3, 96, 96)
    with torch.no_grad():
        want = tv(x)
    plain = EfficientNetB3(weights="imagenet", weights_path=path, include_top=True, include_film=False).eval()
    assert len(plain.state_dict()) == len(tv.state_dict())
    with torch.no_grad():
        got = plain(x)
    torch.testing.assert_close(got, want, rtol=1e-4, atol=1e-5)
    # the FiLM variant takes the same weights by name; its projections are zero-initialised, so it is the same function
    filmed = EfficientNetB3(weights="imagenet", weights_path=path, include_top=True, include_film=True).eval()
    with torch.no_grad():
        got_f = filmed(x, context=torch.randn(1, 512))
    torch.testing.assert_close(got_f, want, rtol=1e-4, atol=1e-5)
    # no-top variant: the feature extractor used by the RT-1 image tokenizer
    path_nt = str(tmp_path / "efficientnetb3_notop.pth")
    torch.save(tv.features.state_dict(), path_nt)
    notop = EfficientNetB3(weights="imagenet", weights_path=path_nt, include_top=False, include_film=False).eval()
    with torch.no_grad():
        torch.testing.assert_close(notop(x), tv.features(x), rtol=1e-4, atol=1e-5)


def test_missing_blob_is_a_clear_error(tmp_path):
    with pytest.raises(FileNotFoundError, match="ImageNet EfficientNet-B3 weights not found"):
        EfficientNetB3(weights="imagenet", weights_path=str(tmp_path / "nope.pth"))


def test_ilsvrc_predictor_names_the_top_classes():
    pred = ILSVRCPredictor(top=3)
    assert len(pred.class_index) == 1000
    out = torch.zeros(1, 1000)
    out[0, 281], out[0, 0], out[0, 999] = 3.0, 2.0, 1.0
    names = pred.predict_topk(out)
    assert len(names) == 3 and "tabby" in names[0].lower() and "tench" in names[1].lower()
    with pytest.raises(AssertionError):
        pred.predict_topk(torch.zeros(2, 1000))

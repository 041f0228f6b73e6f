"""CPU: the C-ABI library builds, loads without a GPU, and exports exactly what include/alexcap.h declares;
host-side argument checking fails loudly (no silent CPU fallback)."""
import ctypes as C
import os
import re
import subprocess

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_prototypes():
    text = open(os.path.join(ROOT, 'include', 'alexcap.h')).read()
    return sorted(set(re.findall(r'^ALEXCAP_API\s+[\w\s\*]+?\b(alexcap_\w+)\s*\(', text, flags=re.M)))


def test_library_builds_and_exports_every_header_symbol():
    from imagecaptioning_b200 import _lib, build
    path = build.build_library()
    names = header_prototypes()
    assert len(names) >= 20
    out = subprocess.run(['nm', '-D', '--defined-only', path], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r' T (alexcap_\w+)', out))
    assert exported == set(names), (sorted(exported - set(names)), sorted(set(names) - exported))
    assert set(_lib.PROTOTYPES) == set(names)            # the ctypes binding covers the whole header
    lib = _lib.load()
    assert lib.alexcap_abi_version() == 2


def test_struct_layout_matches_header():
    from imagecaptioning_b200 import _lib
    # 10 int32 + 3 float... : B,L,D,H,E,Vp,S,num_layers,dtype,need_dfeat,ignore_index (11 x 4) + 3 floats + pad + 2 x u64
    assert C.sizeof(_lib.Dims) == 80      # ABI 2: + uint64_t* philox_state
    assert C.sizeof(_lib.AttnParams) == 19 * 8
    assert C.sizeof(_lib.LstmParams) == (5 + 4 * 8) * 8
    d = _lib.Dims(2, 3, 4, 4, 4, 8, 3, 1, 0, 0, 0, 0.1, 1.0, 0.0, 0, 0)
    lib = _lib.load()
    assert lib.alexcap_attn_workspace_bytes(C.byref(d)) > 0
    assert lib.alexcap_lstm_workspace_bytes(C.byref(d)) > 0
    assert lib.alexcap_attn_decode_workspace_bytes(C.byref(d), 3) > 0


def test_bad_arguments_are_reported_not_swallowed():
    from imagecaptioning_b200 import _lib
    lib = _lib.load()
    d = _lib.Dims(2, 3, 5, 4, 4, 8, 3, 1, 0, 0, 0, 0.1, 1.0, 0.0, 0, 0)      # D=5 not a multiple of 4
    P = _lib.AttnParams()
    rc = lib.alexcap_attn_train_fused(C.byref(d), C.byref(P), None, None, C.byref(P), None, None, None, None, None, None, 0, None)
    assert rc == -1 and b'multiples of 4' in lib.alexcap_last_error()
    d.D = 4
    rc = lib.alexcap_attn_train_fused(C.byref(d), C.byref(P), None, None, C.byref(P), None, None, None, None, None, None, 0, None)
    assert rc == -1 and b'NULL' in lib.alexcap_last_error()
    d.dtype = 7                                                              # neither ALEXCAP_F32 nor ALEXCAP_BF16
    rc = lib.alexcap_attn_forward(C.byref(d), C.byref(P), None, None, None, None, None, None, 0, None)
    assert rc == -2


def test_no_cpu_fallback_in_product_path():
    """CPU tensors must be refused by the modules (the oracle is never reachable from the package)."""
    from imagecaptioning_b200._lib import AlexcapError
    from imagecaptioning_b200.modules import Decoder
    dec = Decoder(10, 8, 8, 8, 4, {}, 0.0, 'cpu')
    with pytest.raises((AlexcapError, RuntimeError)):
        dec(torch.randn(2, 3, 8), torch.ones(2, 4, dtype=torch.int32))
    src = ''
    pkg = os.path.join(ROOT, 'imagecaptioning_b200')
    for f in os.listdir(pkg):
        if f.endswith('.py'):
            src += open(os.path.join(pkg, f)).read()
    assert 'import oracle' not in src and 'from oracle' not in src


def test_state_dict_names_match_reference():
    from imagecaptioning_b200.modules import Decoder, LanguageModel
    dec = Decoder(50, 16, 32, 24, 5, {}, 0.0, 'cpu')
    assert list(dec.state_dict().keys()) == [
        'init_h.weight', 'init_h.bias', 'init_c.weight', 'init_c.bias', 'f_beta.weight', 'f_beta.bias',
        'deep_output.weight', 'deep_output.bias', 'attention.U.weight', 'attention.U.bias', 'attention.W.weight',
        'attention.W.bias', 'attention.v.weight', 'attention.v.bias', 'embedding.weight', 'lstm.weight_ih',
        'lstm.weight_hh', 'lstm.bias_ih', 'lstm.bias_hh']
    lm = LanguageModel(50, 16, 32, 24, 5, 1, {}, 0.0, 'cpu')
    assert set(lm.state_dict().keys()) == {
        'image_encoder.encode.weight', 'image_encoder.encode.bias', 'lookup_table.weight', 'rnn.linear.weight',
        'rnn.linear.bias', 'lstm.weight_ih_l0', 'lstm.weight_hh_l0', 'lstm.bias_ih_l0', 'lstm.bias_hh_l0'}
    assert dec.dropout.p == 0.5 and dec.use_beam and dec.beam_size == 5 and dec.START_TOKEN == 51 and dec.END_TOKEN == 52
    assert not lm.use_beam and lm.beam_size == 3


def test_reference_checkpoint_loads(reference_modules):
    from imagecaptioning_b200.modules import Decoder, LanguageModel
    torch.manual_seed(123)
    ref = reference_modules['Decoder'](50, 16, 32, 24, 5, {}, 0.0, 'cpu')
    Decoder(50, 16, 32, 24, 5, {}, 0.0, 'cpu').load_state_dict(ref.state_dict(), strict=True)
    ref = reference_modules['LanguageModel'](50, 16, 32, 24, 5, 2, {}, 0.0, 'cpu')
    LanguageModel(50, 16, 32, 24, 5, 2, {}, 0.0, 'cpu').load_state_dict(ref.state_dict(), strict=True)

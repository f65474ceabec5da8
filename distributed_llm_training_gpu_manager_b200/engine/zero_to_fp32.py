#!/usr/bin/env python
"""Consolidate a ZeRO checkpoint written by this framework into one fp32 ``state_dict`` file —
the counterpart of DeepSpeed's ``zero_to_fp32.py`` (which it copies next to every checkpoint; the
reference promises those files but never writes any). This copy understands the
row-partition layout recorded under ``b200zero_layout`` in the shard files and needs only torch.

    python zero_to_fp32.py <checkpoint_dir> <output.pt> [--tag global_step100] [--hf_config config.json]

With ``--hf_config`` (a transformers Llama ``config.json``) the tensors are written under the transformers names
(``q|k|v`` and ``gate|up`` split again, vocabulary padding dropped), ready for ``LlamaForCausalLM.load_state_dict``.
"""
import argparse
import json
import sys
from pathlib import Path

import torch


def consolidate(ckpt_dir: str, tag: str = None):
    root = Path(ckpt_dir)
    tag = tag or (root / "latest").read_text().strip()
    tdir = root / tag
    manifest = json.loads((tdir / "manifest.json").read_text())
    ws = manifest["world_size"]
    shards = []
    for r in range(ws):
        hits = sorted(tdir.glob(f"*zero_pp_rank_{r}_mp_rank_00_optim_states.pt"))
        if not hits:
            raise FileNotFoundError(f"optimizer shard of rank {r} missing in {tdir}")
        shards.append(torch.load(hits[0], map_location="cpu", weights_only=False))
    layout = shards[0]["b200zero_layout"]
    flats = []
    for sd in shards:
        osd = sd["optimizer_state_dict"]
        flats.append((osd.get("fp32_flat_groups") or osd.get("single_partition_of_fp32_groups"))[0])
    out = {}
    for unit in layout["units"]:
        for p in unit["params"]:
            off, n, shape = p["shard_offset"], p["shard_numel"], tuple(p["shape"])
            if p["partitioned"]:
                rows = [flats[r][off:off + n].view((shape[0] // ws,) + shape[1:]) for r in range(ws)]
                out[p["name"]] = torch.cat(rows, 0).clone()
            else:
                out[p["name"]] = flats[0][off:off + n].view(shape).clone()
    return out, tag


def to_hf_llama(sd, hf_cfg: dict):
    """This framework's names -> transformers Llama names (same convention as ``models/hf.py``, restated here so
    the script keeps working when copied next to a checkpoint)."""
    d, L = hf_cfg["hidden_size"], hf_cfg["num_hidden_layers"]
    heads = hf_cfg["num_attention_heads"]
    kv = hf_cfg.get("num_key_value_heads") or heads
    hd = hf_cfg.get("head_dim") or d // heads
    nq, nk, f, v = heads * hd, kv * hd, hf_cfg["intermediate_size"], hf_cfg["vocab_size"]
    out = {"model.embed_tokens.weight": sd["embed.weight"][:v].clone(), "model.norm.weight": sd["final_norm.weight"].clone()}
    if not hf_cfg.get("tie_word_embeddings", False):
        out["lm_head.weight"] = sd["lm_head.weight"][:v].clone()
    for i in range(L):
        h, p = f"model.layers.{i}.", f"blocks.{i}."
        qkv, gu = sd[p + "wqkv.weight"], sd[p + "w_gate_up.weight"]
        out[h + "self_attn.q_proj.weight"], out[h + "self_attn.k_proj.weight"] = qkv[:nq].clone(), qkv[nq:nq + nk].clone()
        out[h + "self_attn.v_proj.weight"] = qkv[nq + nk:].clone()
        out[h + "self_attn.o_proj.weight"] = sd[p + "wo.weight"].clone()
        out[h + "input_layernorm.weight"] = sd[p + "attn_norm.weight"].clone()
        out[h + "post_attention_layernorm.weight"] = sd[p + "ffn_norm.weight"].clone()
        out[h + "mlp.gate_proj.weight"], out[h + "mlp.up_proj.weight"] = gu[:f].clone(), gu[f:].clone()
        out[h + "mlp.down_proj.weight"] = sd[p + "w_down.weight"].clone()
    return out


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("checkpoint_dir")
    ap.add_argument("output_file")
    ap.add_argument("--tag", default=None)
    ap.add_argument("--hf_config", default=None, help="transformers Llama config.json: write transformers parameter names")
    ns = ap.parse_args(argv)
    sd, tag = consolidate(ns.checkpoint_dir, ns.tag)
    if ns.hf_config:
        sd = to_hf_llama(sd, json.loads(Path(ns.hf_config).read_text()))
    torch.save(sd, ns.output_file)
    print(f"{tag}: wrote {len(sd)} tensors ({sum(t.numel() for t in sd.values()) / 1e6:.1f} M parameters) to {ns.output_file}")
    return 0


if __name__ == "__main__":
    sys.exit(main())

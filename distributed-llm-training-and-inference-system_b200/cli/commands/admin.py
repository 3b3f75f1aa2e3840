"""``llmctl admin gc | inspect | index`` — real implementations of the three reference stubs
(reference llmctl/cli/commands/admin.py:9-29)."""
from __future__ import annotations

import json
from pathlib import Path
from typing import Optional

import typer
from rich.table import Table

from ..common import console, fail, require

app = typer.Typer(help="Administrative utilities")


@app.command("gc")
def gc(
    output_dir: Path = typer.Option(Path("./outputs"), "--output-dir", "--path", help="Run directory holding checkpoint-*/"),
    keep_latest: int = typer.Option(3, "--keep-latest", help="Complete checkpoints to keep"),
    dry_run: bool = typer.Option(False, "--dry-run"),
):
    """Delete old / incomplete checkpoints (the ``keep_latest`` policy of the training preset)."""
    from ...io.checkpoint import gc_checkpoints
    doomed = gc_checkpoints(output_dir, keep_latest, dry_run=dry_run)
    for d in doomed:
        console.print(("would delete " if dry_run else "deleted ") + str(d))
    console.print(f"[green]{len(doomed)} checkpoint(s) {'would be ' if dry_run else ''}removed[/green]")


@app.command("inspect")
def inspect(
    checkpoint: Optional[Path] = typer.Option(None, "--checkpoint", "--ckpt", help="Checkpoint directory"),
    stats: bool = typer.Option(False, "--stats", help="Also compute per-tensor mean / std / absmax"),
    limit: int = typer.Option(40, "--limit", help="Rows to print"),
):
    """Show a checkpoint's training state and tensor inventory."""
    from safetensors import safe_open
    ck = Path(require(checkpoint, "--checkpoint"))
    if not ck.exists():
        fail(f"checkpoint not found: {ck}", 2)
    ts = ck / "training_state.json"
    if ts.exists():
        st = json.loads(ts.read_text())
        console.print(f"step {st.get('global_step')}  epoch {st.get('epoch')}  mesh {st.get('mesh')}  data {st.get('data')}")
    console.print(f"complete: {(ck / 'COMPLETE').exists()}   optimizer shards: {len(list((ck / 'optim').glob('*.pt'))) if (ck / 'optim').exists() else 0}")
    t = Table(title=str(ck))
    for c in ("file", "tensor", "shape", "dtype") + (("mean", "std", "absmax") if stats else ()):
        t.add_column(c)
    total, rows = 0, 0
    for f in sorted(ck.glob("*.safetensors")):
        with safe_open(str(f), framework="pt") as sf:
            for k in sf.keys():
                tensor = sf.get_tensor(k) if stats or True else None
                total += tensor.numel()
                if rows < limit:
                    row = [f.name, k, str(tuple(tensor.shape)), str(tensor.dtype).replace("torch.", "")]
                    if stats:
                        x = tensor.float()
                        row += [f"{x.mean():.4g}", f"{x.std():.4g}", f"{x.abs().max():.4g}"]
                    t.add_row(*row)
                    rows += 1
    for f in sorted(ck.glob("*.gguf")):                    # llama.cpp containers written by `export --format gguf`
        from ...io.gguf import read_gguf
        meta, tensors = read_gguf(f)
        console.print(f"{f.name}: GGUF v3, architecture {meta.get('general.architecture')}, "
                      f"{meta.get('llama.block_count')} blocks, hidden {meta.get('llama.embedding_length')}, "
                      f"vocabulary {meta.get('llama.vocab_size')}")
        for k, tensor in tensors.items():
            total += tensor.numel()
            if rows < limit:
                row = [f.name, k, str(tuple(tensor.shape)), str(tensor.dtype).replace("torch.", "")]
                if stats:
                    x = tensor.float()
                    row += [f"{x.mean():.4g}", f"{x.std():.4g}", f"{x.abs().max():.4g}"]
                t.add_row(*row)
                rows += 1
    exp = ck / "export.json"
    if exp.exists():
        info = json.loads(exp.read_text())
        console.print(f"export: format {info.get('format')}, quantization {info.get('quantization') or 'none'}")
    console.print(t)
    console.print(f"[green]{total:,} parameters in weight files[/green]")
    return total


@app.command("index")
def index(
    dataset: Optional[Path] = typer.Option(None, "--dataset", help="Text (.txt) or JSON-lines (.jsonl, field 'text') file"),
    out: Optional[Path] = typer.Option(None, "--out", help="Token file to write (default: <dataset>.bin)"),
    vocab_size: int = typer.Option(32000, "--vocab-size"),
    tokenizer: Optional[str] = typer.Option(None, "--tokenizer", help="HF tokenizer directory (default: byte-level)"),
):
    """Tokenise a dataset into a memory-mappable token file + ``.meta.json`` index."""
    from ...io.dataset import build_token_file
    from ...serve.tokenizer import load_tokenizer
    ds = Path(require(dataset, "--dataset"))
    if not ds.exists():
        fail(f"dataset not found: {ds}", 2)
    dst = Path(out) if out else ds.with_suffix(".bin")
    tok = load_tokenizer(tokenizer, vocab_size)
    meta = build_token_file(ds, dst, tok, dtype="uint16" if tok.vocab_size <= 65536 else "uint32")
    console.print(f"[green]Indexed {meta['num_tokens']:,} tokens -> {dst}[/green]")
    return meta

"""``ts2d -i <in> -o <out>`` command line (reference ts2d/main.py:10-115: same flags and file enumeration)."""
import os
import shutil
from glob import glob

from .core.util.config import get_default_model
from .core.util.log import log, log_silent
from .tool import TS2D


def _enumerate_cases(src: str):
    isdir = os.path.isdir(src)
    files = sorted(glob(os.path.join(src, "*.*"))) if isdir else [src]
    for fp in files:
        try:
            if not os.path.exists(fp):
                raise FileNotFoundError(f"Source file does not exist: {fp}")
            if not os.path.isfile(fp):
                raise ValueError(f"Source is not a regular file: {fp}")
            fn = os.path.basename(fp)
            if "." not in fn:
                raise ValueError(f"Source file does not have an extension: {fn}")
            name, ext = fn.split(".", maxsplit=1)
            if ext not in ("nrrd", "nii", "nii.gz", "mha", "mhd"):
                raise ValueError(f"Unsupported file extension: {ext} in {fn}")
            yield name, fp
        except Exception:
            if isdir:
                continue
            raise


def ts2d_run(src: str, dest: str, model: str = None, use_remote: bool = True, fetch_remote: bool = True, collapse: bool = False,
             visualize: bool = True, save_all: bool = False, silent: bool = False):
    model = get_default_model() if model is None else model
    content = "all" if visualize else "file"
    models = "all" if save_all else "final"
    log_silent(silent)
    bar = "#" * shutil.get_terminal_size(fallback=(120, 20)).columns
    log(f"\n{bar}\nTS2D is a research tool. It is NOT validated for clinical use and should NOT be used for medical diagnosis or treatment.\n"
        "This is the B200-native implementation of the TS2D inference path; please cite the TS2D paper when using it:\n"
        "Sabrowsky-Hirsch, B., Alshenoudy, A., Thumfart, S., & Giretzlehner, M. (2025). TotalSegmentator 2D: A Tool for Rapid "
        f"Anatomical Structure Analysis. MIUA 2025, pp. 32-43.\n{bar}\n")
    from . import shard
    rank, world, local = shard.rank_world()
    if world > 1:      # launched under torchrun: one process per GPU, cases sharded round robin, no collective on the data path
        import torch
        torch.cuda.set_device(local)
    with TS2D(key=model, use_remote=use_remote, fetch_remote=fetch_remote) as ts:
        all_cases = list(_enumerate_cases(src))
        cases = shard.shard(all_cases, rank, world)
        n_cases = len(cases)
        log(f"Predicting {n_cases} case{'s' if n_cases != 1 else ''}" + (f" (rank {rank} of {world}, {len(all_cases)} in total)" if world > 1 else ""))
        # pipelined: while case i computes, case i+1 is uploaded and case i-1 is downloaded and written
        results = ts.predict_batch((path for _, path in cases), collapse=collapse)
        for case_id, ((name, path), res) in enumerate(zip(cases, results)):
            log(f"[{case_id + 1}/{n_cases}] Processed: {name}")
            res.save(dest=dest, name=name, models=models, content=content, targets=["segmentation", "projection"])


def ts2d_entry_point():
    import argparse
    parser = argparse.ArgumentParser(description="Runs TotalSegmentator2D (TS2D) on images or directories of images to automatically segment anatomical structures.")
    parser.add_argument("--src", "-i", "--input", type=str, required=True, help="Input image file or directory. Supported formats are: nrrd, nii, nii.gz, mha, mhd")
    parser.add_argument("--dest", "-o", "--output", type=str, required=True, help="Output directory for results.")
    parser.add_argument("--model", type=str, default=None, help="Model key for prediction, defaults to 'ts2d-v2-ep4000b2'.")
    parser.add_argument("--no-remote", action="store_true", help="Disable remote model download. Models must be available locally.")
    parser.add_argument("--no-fetch", action="store_true", help="Do not fetch the latest model URLs from the remote repository.")
    parser.add_argument("--collapse", action="store_true", help="Collapse projected images to 2D. This removes the 3D geometrical information.")
    parser.add_argument("--visualize", action="store_true", help="Visualize the results as PNG images.")
    parser.add_argument("--save-all", action="store_true", help="In addition to the final result, also saves results for each individual model.")
    parser.add_argument("--silent", action="store_true", help="Hides any unnecessary output.")
    args = parser.parse_args()
    ts2d_run(src=args.src, dest=args.dest, model=args.model, use_remote=not args.no_remote, fetch_remote=not args.no_fetch,
             collapse=args.collapse, visualize=args.visualize, save_all=args.save_all, silent=args.silent)


if __name__ == "__main__":
    ts2d_entry_point()

"""Download the raw dataset, as reference ``scripts/data_download.py:7-23``:
``load_dataset(config['dataset_name'], split='train')`` and print a sample.  Needs network access
and the ``datasets`` package; on an offline box it says so instead of raising a stack trace."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from config.config import default_config as config  # noqa: E402


def download_dataset():
    name = config.get("dataset_name", "openwebtext")
    try:
        from datasets import load_dataset
        ds = load_dataset(name, split="train")
    except Exception as e:
        print(f"could not download {name!r}: {type(e).__name__}: {e}\n"
              "(no network / datasets package?  use scripts/data_preprocess.py --text_files ... for local text)")
        return None
    print(f"dataset {name}: {len(ds):,} documents")
    print("sample:", ds[0]["text"][:200])
    return ds


if __name__ == "__main__":
    download_dataset()

def download_cached_file(url, check_hash=True, progress=False):
    raise RuntimeError("no network in this environment: ImageNet_weights=True is unavailable")

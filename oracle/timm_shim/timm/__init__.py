"""Stand-in for the un-installable ``timm==0.6.12`` (reference requirements.txt:7).

TEST INFRASTRUCTURE ONLY.  Restates, from the published timm 0.6.12 behaviour,
just the surface that /root/reference/Models/{models.py,mae/models_mae.py,
moco_v3/vits.py} import, so the *unmodified* reference modules can be imported
in the build container to generate golden vectors (oracle/make_golden.py).
Never imported by the product package ``ssl4gie_b200``.
"""
__version__ = "0.6.12"

"""Allowed MIG geometries per GPU model -- runtime DATA handed to the engine, never compiled into kernels.

DEFAULT_KNOWN_GEOMETRIES reproduces the reference's compiled-in default (pkg/gpu/mig/known_configs.go:24-142) as
given, including the `3g.20gb` / `2g.10gb` entries inside the A100-80GB table (:118,:121).  HELM_KNOWN_GEOMETRIES is
the production table shipped in the chart (helm-charts/nos/values.yaml:317-376); note the different row ORDER, which
matters for the canonical arg-max tie-break (lowest table index among the max scorers).
"""
A30 = "NVIDIA-A30"
A100_40 = "NVIDIA-A100-40GB-SXM4"
A100_80 = "NVIDIA-A100-80GB-PCIe"

DEFAULT_KNOWN_GEOMETRIES = {
    A30: [
        {"4g.24gb": 1}, {"2g.12gb": 2}, {"2g.12gb": 1, "1g.6gb": 2}, {"1g.6gb": 4},
    ],
    A100_40: [
        {"7g.40gb": 1},
        {"4g.20gb": 1, "2g.10gb": 1, "1g.5gb": 1},
        {"4g.20gb": 1, "1g.5gb": 3},
        {"3g.20gb": 2},
        {"3g.20gb": 1, "2g.10gb": 1, "1g.5gb": 1},
        {"3g.20gb": 1, "1g.5gb": 3},
        {"2g.10gb": 2, "3g.20gb": 1},
        {"2g.10gb": 1, "1g.5gb": 2, "3g.20gb": 1},
        {"2g.10gb": 3, "1g.5gb": 1},
        {"2g.10gb": 2, "1g.5gb": 3},
        {"2g.10gb": 1, "1g.5gb": 5},
        {"1g.5gb": 7},
    ],
    A100_80: [
        {"7g.79gb": 1},
        {"4g.40gb": 1, "2g.20gb": 1, "1g.10gb": 1},
        {"4g.40gb": 1, "1g.10gb": 3},
        {"3g.40gb": 2},
        {"3g.40gb": 1, "2g.20gb": 1, "1g.10gb": 1},
        {"3g.40gb": 1, "1g.10gb": 3},
        {"2g.20gb": 2, "3g.20gb": 1},
        {"2g.10gb": 1, "1g.10gb": 2, "3g.40gb": 1},
        {"2g.20gb": 3, "1g.10gb": 1},
        {"2g.20gb": 2, "1g.10gb": 3},
        {"2g.20gb": 1, "1g.10gb": 5},
        {"1g.10gb": 7},
    ],
}

HELM_KNOWN_GEOMETRIES = {
    "A30": [
        {"1g.6gb": 4}, {"1g.6gb": 2, "2g.12gb": 1}, {"2g.12gb": 2}, {"4g.24gb": 1},
    ],
    **{m: [
        {"1g.5gb": 7},
        {"1g.5gb": 5, "2g.10gb": 1},
        {"1g.5gb": 3, "2g.10gb": 2},
        {"1g.5gb": 1, "2g.10gb": 3},
        {"1g.5gb": 2, "2g.10gb": 1, "3g.20gb": 1},
        {"2g.10gb": 2, "3g.20gb": 1},
        {"1g.5gb": 3, "3g.20gb": 1},
        {"1g.5gb": 1, "2g.10gb": 1, "3g.20gb": 1},
        {"3g.20gb": 2},
        {"1g.5gb": 3, "4g.20gb": 1},
        {"1g.5gb": 1, "2g.10gb": 1, "4g.20gb": 1},
        {"7g.40gb": 1},
    ] for m in ("A100-SXM4-40GB", "NVIDIA-A100-40GB-PCIe", "NVIDIA-A100-SXM4-40GB")},
    **{m: [
        {"1g.10gb": 7},
        {"1g.10gb": 5, "2g.20gb": 1},
        {"1g.10gb": 3, "2g.20gb": 2},
        {"1g.10gb": 1, "2g.20gb": 3},
        {"1g.10gb": 2, "2g.20gb": 1, "3g.40gb": 1},
        {"2g.20gb": 2, "3g.20gb": 1},
        {"1g.10gb": 3, "3g.40gb": 1},
        {"1g.10gb": 1, "2g.20gb": 1, "3g.40gb": 1},
        {"3g.40gb": 2},
        {"1g.10gb": 3, "4g.40gb": 1},
        {"1g.10gb": 1, "2g.20gb": 1, "4g.40gb": 1},
        {"7g.79gb": 1},
    ] for m in ("NVIDIA-A100-SXM4-80GB", "NVIDIA-A100-80GB-PCIe")},
}

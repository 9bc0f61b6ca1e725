"""Table files either side of the loader.

    python -m mitsuba_customization_b200.tabletool report  table.binary
    python -m mitsuba_customization_b200.tabletool pack    table.binary table.merlb200 [--linear] [--scale r g b]
    python -m mitsuba_customization_b200.tabletool synth   out.binary [--id 0]

`report` scans a MERL-layout .binary on the host (no GPU needed) and prints what the loader will
see; `pack` loads it on the GPU (prescale, float4 repack, CDF quadrature) and writes the sidecar
that `MERL(packed=...)` / merl_table_create_from_packed upload as is; `synth` writes one of the
synthetic tables the tests and bench use.  All work is done by libmerl_b200 through the C ABI.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import sys

from . import _cabi


def report(path: str) -> dict:
    r = _cabi.FileReport()
    _cabi.check(_cabi.lib().merl_file_report(str(path).encode(), C.byref(r)))
    return {
        "dims": list(r.dims), "is_merl_layout": bool(r.is_merl_layout), "ncell": r.ncell,
        "file_bytes": r.file_bytes, "negative": list(r.negative), "nonfinite": list(r.nonfinite),
        "min_valid": list(r.min_valid), "max_valid": list(r.max_valid),
    }


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="tabletool", description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    sub = ap.add_subparsers(dest="cmd", required=True)
    a = sub.add_parser("report"); a.add_argument("path")
    a = sub.add_parser("pack"); a.add_argument("path"); a.add_argument("out")
    a.add_argument("--linear", action="store_true", help="customized_measurment with linear theta_h bins")
    a.add_argument("--scale", nargs=3, type=float, default=None)
    a.add_argument("--device", default="cuda:0")
    a = sub.add_parser("synth"); a.add_argument("out"); a.add_argument("--id", type=int, default=0)
    args = ap.parse_args(argv)
    try:
        if args.cmd == "report":
            print(json.dumps(report(args.path), indent=1))
        elif args.cmd == "pack":
            from . import bsdf
            rep = report(args.path)
            if rep["is_merl_layout"] and not args.linear and args.scale is None:
                plug = bsdf.MERL(filename=args.path, device=args.device)
            else:
                kw = {"filename": args.path, "device": args.device, "theta_h_mapping": "linear" if args.linear else "sqrt"}
                if args.scale is not None:
                    kw["scale"] = tuple(args.scale)
                plug = bsdf.CustomizedMeasurement(kw)
            plug.save_packed(args.out)
            print(f"{args.out}: {plug.info.nh}x{plug.info.nd}x{plug.info.np}, {plug.info.table_bytes} table bytes")
        else:
            from . import synth
            synth.write_merl_binary(args.out, synth.synthetic_table(args.id))
            print(f"{args.out}: synthetic MERL table id {args.id}")
    except _cabi.MerlError as e:
        print(f"tabletool: {e}", file=sys.stderr)
        return 1
    return 0


if __name__ == "__main__":
    sys.exit(main())

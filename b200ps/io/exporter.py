"""Model export (R13, P26): ``tensorflow.contrib.session_bundle.exporter.Exporter``.

model_export.py:54-66 does::

    exporter.Exporter(saver).init(graph_def, named_graph_signatures={
        'inputs': generic_signature({'x': x}), 'outputs': generic_signature({'y': y_pred})})
    exporter.export('./model/', tf.constant('0003'), sess)

which writes ``<work_dir>/00000003/`` (the version string is parsed as an int and rendered
``%08d``) holding ``export.meta`` (a MetaGraphDef whose ``serving_signatures`` collection
carries the named signatures), the variables as a Saver bundle named ``export`` and an
``assets/`` directory.  The same layout is produced here.  ``export.meta`` is a structurally
valid MetaGraphDef protobuf (saver_def + collection_def with the Signatures message); it has
no executable GraphDef because there is no TF graph — ``signature.json`` states the same
bindings for non-TF consumers (and is what :func:`load_export` reads back).
"""
from __future__ import annotations

import json
import os
import shutil
from typing import Dict, Optional

from . import proto as pb
from . import tf_bundle

VERSION_FORMAT = "%08d"
EXPORT_BASENAME = "export"


def generic_signature(name_tensor_map: Dict[str, str]) -> Dict[str, str]:
    """alias -> tensor name, like ``exporter.generic_signature({'x': x})`` (tensor names are strings here)."""
    return dict(name_tensor_map)


def _signatures_proto(named: Dict[str, Dict[str, str]]) -> bytes:
    # Signatures{ map<string,Signature> named_signatures = 2 } ; Signature{ GenericSignature generic_signature = 3 }
    # GenericSignature{ map<string,TensorBinding> map = 1 } ; TensorBinding{ string tensor_name = 1 }
    out = b""
    for sig_name in sorted(named):
        gs = b""
        for alias in sorted(named[sig_name]):
            binding = pb.f_bytes(1, named[sig_name][alias])
            gs += pb.f_msg(1, pb.f_bytes(1, alias) + pb.f_msg(2, binding))
        signature = pb.f_msg(3, gs)
        out += pb.f_msg(2, pb.f_bytes(1, sig_name) + pb.f_msg(2, signature))
    return out


def _meta_graph_proto(named: Dict[str, Dict[str, str]], variable_names) -> bytes:
    # MetaGraphDef{ MetaInfoDef meta_info_def=1; GraphDef graph_def=2; SaverDef saver_def=3;
    #               map<string,CollectionDef> collection_def=4 }
    meta_info = pb.f_bytes(1, "b200ps_export")                      # meta_graph_version
    saver_def = (pb.f_bytes(1, "save/Const:0") + pb.f_bytes(2, "save/control_dependency:0") +
                 pb.f_bytes(3, "save/restore_all") + pb.f_varint(4, 5) + pb.f_varint(5, 1) +
                 pb.f_varint(7, 2))                                  # sharded, version V2
    any_msg = (pb.f_bytes(1, "type.googleapis.com/tensorflow.serving.Signatures") +
               pb.f_bytes(2, _signatures_proto(named)))
    coll_sig = pb.f_msg(5, pb.f_msg(1, any_msg))                     # CollectionDef.any_list.value[0]
    entries = pb.f_msg(4, pb.f_bytes(1, "serving_signatures") + pb.f_msg(2, coll_sig))
    names = pb.f_msg(2, b"".join(pb.f_bytes(1, n) for n in sorted(variable_names)))   # bytes_list
    entries += pb.f_msg(4, pb.f_bytes(1, "variables") + pb.f_msg(2, names))
    return pb.f_msg(1, meta_info) + pb.f_msg(2, b"") + pb.f_msg(3, saver_def) + entries


class Exporter:
    def __init__(self, saver=None):
        self.saver = saver
        self._signatures: Optional[Dict[str, Dict[str, str]]] = None
        self._assets = None

    def init(self, graph_def=None, named_graph_signatures: Optional[Dict[str, Dict[str, str]]] = None,
             assets: Optional[Dict[str, str]] = None) -> None:
        self._signatures = {k: dict(v) for k, v in (named_graph_signatures or {}).items()}
        self._assets = assets or {}

    def export(self, export_dir_base: str, global_step_tensor, variables: Dict[str, object]) -> str:
        """Write ``export_dir_base/%08d`` and return its path.  ``global_step_tensor`` may be an int
        or the string the reference passes (``tf.constant('0003')`` -> 3)."""
        if self._signatures is None:
            raise RuntimeError("Exporter.init() must be called before export()")
        version = int(global_step_tensor)
        final = os.path.join(export_dir_base, VERSION_FORMAT % version)
        tmp = final + "-tmp"
        if os.path.exists(tmp):
            shutil.rmtree(tmp)
        os.makedirs(os.path.join(tmp, "assets"))
        tf_bundle.write_bundle(os.path.join(tmp, EXPORT_BASENAME), variables)
        with open(os.path.join(tmp, EXPORT_BASENAME + ".meta"), "wb") as f:
            f.write(_meta_graph_proto(self._signatures, variables.keys()))
        with open(os.path.join(tmp, "signature.json"), "w") as f:
            json.dump({"version": version, "named_graph_signatures": self._signatures,
                       "variables": sorted(variables)}, f, indent=1, sort_keys=True)
        for name, src in self._assets.items():
            shutil.copy(src, os.path.join(tmp, "assets", name))
        if os.path.exists(final):
            shutil.rmtree(final)
        os.replace(tmp, final)
        return final


def load_export(export_dir: str):
    """(signatures, variables) of an export directory written by :class:`Exporter`."""
    with open(os.path.join(export_dir, "signature.json")) as f:
        manifest = json.load(f)
    return manifest["named_graph_signatures"], tf_bundle.read_bundle(os.path.join(export_dir, EXPORT_BASENAME))


def parse_signatures_from_meta(path: str) -> Dict[str, Dict[str, str]]:
    """Decode the named signatures back out of ``export.meta`` (checks the protobuf is well formed)."""
    m = pb.decode(open(path, "rb").read())
    for entry in m.get(4, []):
        e = pb.decode(entry)
        if e[1][0] != b"serving_signatures":
            continue
        any_list = pb.decode(pb.decode(e[2][0])[5][0])
        sigs = pb.decode(pb.decode(any_list[1][0])[2][0])
        out = {}
        for named in sigs.get(2, []):
            nm = pb.decode(named)
            generic = pb.decode(pb.decode(nm[2][0])[3][0])
            out[nm[1][0].decode()] = {
                pb.decode(kv)[1][0].decode(): pb.decode(pb.decode(kv)[2][0])[1][0].decode()
                for kv in generic.get(1, [])}
        return out
    return {}

"""Oracle (test infrastructure): the pure part of the trusted-CA bundle path.

Reference: components/odh-notebook-controller/controllers/notebook_controller.go
  CreateNotebookCertConfigMap :254-357, IsConfigMapDeleted :361-387,
  UnsetNotebookCertConfig :390-455.
The ConfigMap reads and writes are I/O and stay with the caller; what is restated here is
what the reference computes from the bytes it read: which of the three certificate files
make it into the bundle (first PEM block of each must be a CERTIFICATE that parses), the
joined bundle, and the Notebook with the certificate configuration taken out again —
including the delete-while-ranging loops of :411-415 and :418-422, simulated on Go's slice
semantics (a range over a slice keeps the original length and reads the shared backing array).

Go's encoding/pem and crypto/x509 are not restated in full: `pem_decode` follows pem.Decode
for the shapes certificates come in (BEGIN/END lines, optional headers, base64 body), and
x509.ParseCertificate is approximated by `der_is_certificate` (the outer DER structure of a
Certificate); callers with a real X.509 parser pass its verdicts in `x509_ok`.
"""
import base64
import binascii
import re
from copy import deepcopy

ENV_VARS = ["PIP_CERT", "REQUESTS_CA_BUNDLE", "SSL_CERT_FILE", "PIPELINES_SSL_SA_CERTS", "GIT_SSL_CAINFO"]   # :396
GO_SPACE = " \t\n\v\f\r\x85\xa0"


class ReferencePanic(Exception):
    """The reference would panic here (slice bounds out of range)."""


def trim_space(s):
    return s.strip(GO_SPACE)                                                            # strings.TrimSpace :289


def pem_decode(data):
    """pem.Decode: (type, der bytes, rest) of the first block, or (None, None, data)."""
    m = re.search(r"(?:^|\n)-----BEGIN ([^\n-]*)-----[ \t\r]*\n", data)
    if not m:
        return None, None, data
    typ = m.group(1)
    body_start = m.end()
    end = re.compile(r"(?:^|\n)-----END " + re.escape(typ) + r"-----[ \t\r]*(?:\n|$)")
    e = end.search(data, body_start - 1)
    if not e:
        return None, None, data
    body = data[body_start:e.start() + (1 if data[e.start()] == "\n" else 0)]
    # optional headers: "Key: value" lines up to a blank line
    lines = body.split("\n")
    while lines and ":" in lines[0]:
        lines.pop(0)
    if lines and lines[0].strip() == "" and len(lines) != len(body.split("\n")):
        lines.pop(0)
    try:
        der = base64.b64decode("".join(x.strip() for x in lines), validate=True)
    except (binascii.Error, ValueError):
        return None, None, data
    return typ, der, data[e.end():]


def _der_tlv(b, i):
    """(tag, content start, content end) of the DER element at i, or None."""
    if i + 2 > len(b):
        return None
    tag, ln = b[i], b[i + 1]
    i += 2
    if ln & 0x80:
        k = ln & 0x7F
        if k == 0 or k > 4 or i + k > len(b):
            return None
        ln = int.from_bytes(b[i:i + k], "big")
        i += k
    if i + ln > len(b):
        return None
    return tag, i, i + ln


def der_is_certificate(der):
    """Certificate ::= SEQUENCE { tbsCertificate SEQUENCE, signatureAlgorithm SEQUENCE,
    signatureValue BIT STRING }, lengths exact.  A structural stand-in for
    x509.ParseCertificate (:304)."""
    outer = _der_tlv(der, 0)
    if not outer or outer[0] != 0x30 or outer[2] != len(der):
        return False
    i = outer[1]
    for want in (0x30, 0x30, 0x03):
        t = _der_tlv(der, i)
        if not t or t[0] != want or t[2] > outer[2]:
            return False
        i = t[2]
    return i == outer[2]


def cert_bundle(odh_cm, kube_cm, x509_ok=None):
    """What CreateNotebookCertConfigMap puts into workbench-trusted-ca-bundle["ca-bundle.crt"],
    or None when it writes nothing.  odh_cm / kube_cm: the .data dicts of odh-trusted-ca-bundle
    and kube-root-ca.crt, None = NotFound.  x509_ok: optional {file name: bool} verdicts of a
    real x509 parser for the first block of each file."""
    pool = []
    for cm, files, is_odh in ((odh_cm, ["ca-bundle.crt", "odh-ca-bundle.crt"], True), (kube_cm, ["ca.crt"], False)):
        if cm is None:
            if is_odh:
                return None                                                             # :277-279
            continue                                                                    # :280-281
        for f in files:
            ok = f in cm
            data = trim_space(cm.get(f, ""))
            if not ok or (f == "ca-bundle.crt" and data == ""):                         # :293-295
                return None
            if data == "":                                                              # :296-298
                continue
            typ, der, _ = pem_decode(data)                                              # :301
            if typ == "CERTIFICATE":
                good = x509_ok[f] if x509_ok is not None and f in x509_ok else der_is_certificate(der)
                if good:
                    pool.append(data)                                                   # :310 the whole file
    if not pool:
        return None
    return "\n".join(pool)                                                              # :325


def trusted_ca_configmap(namespace, bundle):
    return {"apiVersion": "v1", "kind": "ConfigMap",
            "metadata": {"name": "workbench-trusted-ca-bundle", "namespace": namespace,
                         "labels": {"opendatahub.io/managed-by": "workbenches"}},
            "data": {"ca-bundle.crt": bundle}}                                          # :317-327


def count_certificates(bundle):
    """checkCertConfigMap of the reference's test (notebook_controller_test.go:1134-1159)."""
    n, rest = 0, bundle
    while rest:
        typ, der, rest = pem_decode(rest)
        if typ is None:
            break
        if typ == "CERTIFICATE":
            assert der_is_certificate(der)
            n += 1
    return n


def is_configmap_deleted(nb, workbench_cm_exists):
    """IsConfigMapDeleted :361-387."""
    if workbench_cm_exists:
        return False
    vols = (((nb.get("spec") or {}).get("template") or {}).get("spec") or {}).get("volumes") or []
    return any(isinstance(v.get("configMap"), dict) and v["configMap"].get("name") == "workbench-trusted-ca-bundle"
               for v in vols)


def _delete_while_ranging(items, match):
    """`for index, x := range s { if match(x) { s = append(s[:index], s[index+1:]...) } }`:
    the range keeps the original length and reads the backing array the appends shift."""
    backing = list(items)
    length = len(backing)
    for index in range(len(items)):           # the range expression is evaluated once
        x = backing[index]
        if match(x):
            if index + 1 > length:            # s[index+1:] beyond len(s)
                raise ReferencePanic("slice bounds out of range [%d:%d]" % (index + 1, length))
            tail = backing[index + 1:length]
            backing[index:index + len(tail)] = tail
            length -= 1
    return backing[:length]


def unset_notebook_cert_config(nb):
    """UnsetNotebookCertConfig :390-455: (mutated Notebook, changed)."""
    out = deepcopy(nb)
    name = (out.get("metadata") or {}).get("name", "")
    spec = ((out.get("spec") or {}).get("template") or {}).get("spec") or {}
    changed = False
    ctrs = spec.get("containers") or []
    for c in ctrs:
        if c.get("name", "") == name:                                                   # :408
            img = dict(c)
            env = img.get("env")
            for key in ENV_VARS:                                                        # :410-416
                if env:
                    env = _delete_while_ranging(env, lambda e, key=key: e.get("name", "") == key)
            if env is not None:
                img["env"] = env
            vm = img.get("volumeMounts")
            if vm:
                img["volumeMounts"] = _delete_while_ranging(vm, lambda m: m.get("name", "") == "trusted-ca")   # :418-422
            for i, c2 in enumerate(ctrs):                                               # :424-430
                if c2.get("name", "") == name:
                    ctrs[i] = img
                    changed = True
                    break
            break
    vols = spec.get("volumes") or []
    for i, v in enumerate(vols):                                                        # :436-442
        if isinstance(v.get("configMap"), dict) and v["configMap"].get("name") == "workbench-trusted-ca-bundle":
            del vols[i]
            changed = True
            break
    return out, changed

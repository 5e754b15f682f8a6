"""The INTEGRATION.md patch, compiled: oracle/_ref/refgzip_b200 is the reference's own src/zip.c with
integration/zip_b200.patch applied (zip() -> gzb_zip_fd()), linked with the reference's unmodified parallel.c / unzip.c
and libgzb200.so (oracle/Makefile).  This runs the reference's tests/parallel matrix through it -- -j {1,2,4,8} x
{redirect from a file, pipe} x {a large text, the 9-byte "parallel\\n"} -- and decodes with the UNPATCHED reference
`gzip -d` (oracle/_ref/refgzip -d) and the system gunzip.  (tests/parallel's third mode, `gzip -j N FILE`, is src/gzip.c's
file handling, which is outside the path and outside the shim.)"""
import os, subprocess
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PATCHED = os.path.join(ROOT, "oracle", "_ref", "refgzip_b200")
STOCK = os.path.join(ROOT, "oracle", "_ref", "refgzip")


@pytest.fixture(scope="module")
def inputs(gz, tmp_path_factory):
    if not (os.path.exists(PATCHED) and os.path.exists(STOCK)):
        pytest.skip("oracle/_ref/refgzip_b200 is built where /root/reference exists (python __graft_entry__.py)")
    d = tmp_path_factory.mktemp("par")
    cases = {"large": gz.corpus("text", (2 << 20) + 4567, seed=19).tobytes(),        # tests/parallel: `configure`, a large text file
             "pipeline": gz.corpus("mixed", (40 << 20) + 99, seed=20).tobytes(),      # above gzb_zip_fd's small-input threshold: reader / workers / writer
             "small": b"parallel\n"}
    paths = {}
    for k, v in cases.items():
        p = d / k; p.write_bytes(v); paths[k] = str(p)
    return cases, paths


@pytest.mark.parametrize("j", [1, 2, 4, 8])
@pytest.mark.parametrize("mode", ["redirect", "pipe"])
def test_reference_parallel_matrix_through_the_patched_zip(inputs, j, mode):
    cases, paths = inputs
    for name, data in cases.items():
        if name == "pipeline" and (j not in (1, 8)):
            continue
        if mode == "redirect":
            with open(paths[name], "rb") as fi:
                p = subprocess.run([PATCHED, "-j", str(j)], stdin=fi, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        else:
            cat = subprocess.Popen(["cat", paths[name]], stdout=subprocess.PIPE)
            p = subprocess.run([PATCHED, "-j", str(j)], stdin=cat.stdout, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
            cat.wait()
        assert p.returncode == 0, p.stderr
        member = p.stdout
        assert member[:4] == b"\x1f\x8b\x08\x00" and member[8:10] == b"\x00\x03"          # no FNAME from stdin, XFL 0 at -6, OS = Unix
        back = subprocess.run([STOCK, "-d"], input=member, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        assert back.returncode == 0 and back.stdout == data, (name, j, mode)
        assert subprocess.run(["gunzip", "-c"], input=member, stdout=subprocess.PIPE, check=True).stdout == data


def test_patched_zip_levels_rsyncable_and_sizes(inputs, golden):
    cases, paths = inputs
    data = cases["large"]
    for args in (["-1"], ["-9"], ["-6", "--rsyncable"]):
        with open(paths["large"], "rb") as fi:
            p = subprocess.run([PATCHED, *args], stdin=fi, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        assert p.returncode == 0, p.stderr
        assert subprocess.run([STOCK, "-d"], input=p.stdout, stdout=subprocess.PIPE, check=True).stdout == data
        with open(paths["large"], "rb") as fi:
            ref = subprocess.run([STOCK, *args, "-j", "4"], stdin=fi, stdout=subprocess.PIPE, check=True).stdout
        assert p.stdout[:10] == ref[:10] and p.stdout[-8:] == ref[-8:]                       # header, CRC32 | ISIZE: byte-equal
        assert len(p.stdout) <= 1.03 * len(ref)


def test_reference_boundary_sizes_through_the_patched_zip(gz, inputs, tmp_path):
    """The size list of the reference's tests/mixed:51-57 (0..64 sampled, 32831-32833, 131071-131073 -- the edges of its
    window and of a chunk) compressed by the patched binary from a regular file and from a pipe, decoded by the unpatched
    reference.  Size 0 gives the valid 20-byte member (the reference's own -j path deadlocks there)."""
    for n in (0, 1, 2, 3, 31, 32, 33, 63, 64, 32831, 32832, 32833, 131071, 131072, 131073):
        data = gz.corpus("zipf", n, seed=n).tobytes()
        f = tmp_path / "in"; f.write_bytes(data)
        with open(f, "rb") as fi:
            a = subprocess.run([PATCHED, "-6", "-j", "2"], stdin=fi, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        assert a.returncode == 0, (n, a.stderr)
        if n in (0, 64, 131073):                                 # every process start is a CUDA context: the pipe variant for three of them
            b = subprocess.run([PATCHED, "-6"], input=data, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
            assert b.returncode == 0 and a.stdout == b.stdout, (n, b.stderr)
        if n == 0:
            assert len(a.stdout) == 20
        back = subprocess.run([STOCK, "-d"], input=a.stdout, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        assert back.returncode == 0 and back.stdout == data, n


"""File-read smoke test — the reference's hdfs_test.py: N threads each count the lines of one
input file through the gfile stand-in (hdfs_test.py:4-28).  Pass paths with ``--files a,b,c``
(``hdfs://`` URIs resolve through $B200PS_HDFS_ROOT); with no flag, three synthetic shards."""
from __future__ import annotations

import os
import tempfile
import threading

from .. import flags
from ..data.libsvm import count_lines

FLAGS = flags.FLAGS
flags.DEFINE_string("files", "", "comma-separated files to count")


class LoadDataThread(threading.Thread):
    def __init__(self, tid, file_name):
        super().__init__()
        self.tid, self.file_name, self.lines = tid, file_name, 0

    def run(self):
        print("thread: %d, input file: %s" % (self.tid, self.file_name))
        self.lines = count_lines(self.file_name)
        print("thread: %d, lines: %d" % (self.tid, self.lines))


def main(argv=None) -> int:
    files = [f for f in FLAGS.files.split(",") if f]
    if not files:
        from ..data.synthetic import write_libsvm_files
        files = write_libsvm_files(os.path.join(tempfile.gettempdir(), f"b200ps_hdfs_test_{os.getuid()}"), 3, 100)
    threads = [LoadDataThread(i, f) for i, f in enumerate(files)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    print("total lines:", sum(t.lines for t in threads))
    return 0


if __name__ == "__main__":
    flags.run(main)

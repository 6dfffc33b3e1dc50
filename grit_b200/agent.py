"""Host-side mirror of the two agent actions that call the data mover.

  RunCheckpoint <- pkg/gritagent/checkpoint/checkpoint.go:13-21
  RunRestore    <- pkg/gritagent/restore/restore.go:14-21
  GritAgentOptions <- cmd/grit-agent/app/options/options.go:12-21

The CRI/containerd half of RunCheckpoint (RuntimeCheckpointPod,
checkpoint/runtime.go) is out of scope: it produces the directory this path
consumes.  `python -m grit_b200.agent --action ... --src-dir ... --dst-dir ...`
takes the same flags as cmd/grit-agent for the two directories and the action.
"""
import argparse
import dataclasses
import os
import sys

from . import copy
from ._lib import GritError

# pkg/metadata/metadata.go:8-9
ContainerLogFile = "container.log"
DownloadSentinelFile = "download-state"

ActionCheckpoint = "checkpoint"  # options.go:33
ActionRestore = "restore"  # options.go:34


@dataclasses.dataclass
class GritAgentOptions:
    Action: str = ""
    SrcDir: str = ""
    DstDir: str = ""
    Device: int = 0
    ParentDir: str = ""  # previous checkpoint on the PV, for an incremental upload
    Verify: bool = True


def RunCheckpoint(opts):
    """checkpoint.go:20 -- transfer checkpointed data to cloud storage."""
    return copy.TransferData(opts.SrcDir, opts.DstDir, copy.UPLOAD, device=opts.Device,
                             parent_dir=opts.ParentDir or None)


def RunRestore(opts):
    """restore.go:16-20 -- download, then the sentinel patched containerd waits for."""
    st = copy.TransferData(opts.SrcDir, opts.DstDir, copy.DOWNLOAD, device=opts.Device, verify=opts.Verify)
    copy.CreateSentinelFile(opts.DstDir, DownloadSentinelFile)
    return st


def Run(opts):
    """cmd/grit-agent/app/app.go:53-72."""
    if opts.Action == ActionCheckpoint:
        return RunCheckpoint(opts)
    if opts.Action == ActionRestore:
        return RunRestore(opts)
    raise ValueError(f"unknown action {opts.Action}")


def main(argv=None):
    ap = argparse.ArgumentParser(prog="grit-agent")
    ap.add_argument("--action", default=os.environ.get("ACTION", ""))
    ap.add_argument("--src-dir", default="")
    ap.add_argument("--dst-dir", default="")
    ap.add_argument("--device", type=int, default=0)
    ap.add_argument("--parent-dir", default="")
    a = ap.parse_args(argv)
    try:
        st = Run(GritAgentOptions(a.action, a.src_dir, a.dst_dir, a.device, a.parent_dir))
    except (GritError, ValueError) as e:
        print(f"run grit-agent failed: {e}", file=sys.stderr)  # app.go:40
        return 1
    print(st)
    return 0


if __name__ == "__main__":
    sys.exit(main())

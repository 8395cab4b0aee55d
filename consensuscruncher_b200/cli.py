"""`ConsensusCruncher.py consensus` on the GPU path - same flags, same output tree.

Mirrors the reference's driver (`ConsensusCruncher.py:128-322`, argparse `:433-443`, config
handling `:386-417,453-455,478-479`): `-i/--input BAM`, `-o/--output DIR`, `-s/--samtools`
(accepted and ignored: sort / merge / index are done in-process),
`--scorrect True|False`, `-g/--genome hg19|hg38`, `-b/--bedfile FILE|False`, `--cutoff F`,
`-d/--bdelim`, `--cleanup True|False`, `-c/--config INI`.  Booleans are the strings
'True' / 'False', as in the reference.

The reference runs three stage scripts that hand sorted BAMs to each other; here one pass
over the packed reads produces every family-level result on the GPU and `emit.py` lays the
records out into the same files (App. A.6):

  OUT/ID/sscs/     ID.sscs.sorted.bam  ID.singleton.sorted.bam  ID.badReads.bam
  OUT/ID/dcs/      ID.dcs.sorted.bam   ID.sscs.singleton.sorted.bam
  OUT/ID/sscs_sc/  ID.sscs.correction.sorted.bam  ID.singleton.correction.sorted.bam
                   ID.uncorrected.sorted.bam  ID.sscs.sc.sorted.bam
  OUT/ID/dcs_sc/   ID.dcs.sc.sorted.bam  ID.sscs.sc.singleton.sorted.bam  ID.all.unique.dcs.sorted.bam
  OUT/ID/          ID.stats.txt  ID.time_tracker.txt  ID.read_families.txt
(+ .bai next to every sorted BAM; the tag-family-size PNG is out of scope.)
"""
from __future__ import annotations

import argparse
import configparser
import os
import sys
import time

from . import bam, emit, fastpack, keys, pack

HERE = os.path.dirname(os.path.abspath(__file__))


def default_bedfile(genome):
    """The cytoband table the reference defaults `-b` to (`ConsensusCruncher.py:382-384`, hg38 forced
    by `-g hg38`, `:143-145`): UCSC's public cytoBand tables, 863 contiguous bands each, shipped as
    data in consensuscruncher_b200/data/.  CC_CYTOBAND_DIR names another directory."""
    name = "%s_cytoBand.txt" % genome
    for d in (os.environ.get("CC_CYTOBAND_DIR"), os.path.join(HERE, "data")):
        if d and os.path.exists(os.path.join(d, name)):
            return os.path.join(d, name)
    return None


def regions_from_bed(bedfile, header):
    """[(tid, start, end)] in bed order; a bed contig missing from the BAM header is an error,
    as `fetch` raises in the reference."""
    out = []
    for key, (start, end) in keys.bed_separator(bedfile).items():
        chrom = key.split("_", 1)[0]
        tid = header.tid(chrom)
        if tid < 0:
            raise ValueError("bed contig %r is not in the BAM header" % chrom)
        out.append((tid, start, end))
    return out


def run_consensus(bam_path, out_dir, cutoff=0.7, scorrect=True, bedfile=None, bdelim="|", cleanup=False,
                  device=0, engine=None, log=print, gpus=1, chunk_reads=20_000_000):
    """Everything `consensus()` does in the reference.  Returns the sample directory.

    The packed reads go through the GPU(s) in chunks of consecutive bed regions of about `chunk_reads`
    reads (stream.py): `gpus` contexts, one per device starting at `device`, take the chunks in turn; the
    output does not depend on either number."""
    from . import stream
    from .engine import GpuConsensus
    t0 = time.time()
    identifier = os.path.basename(bam_path).split(".bam", 1)[0]
    sample_dir = os.path.join(out_dir, identifier)
    if os.path.exists(sample_dir):                         # like the reference, a re-run needs a clean directory (Q19) -
        raise FileExistsError(sample_dir)                  # but nothing is created before the input has been read
    header, reads = fastpack.open_bam(bam_path)            # native inflate + pack; records decode lazily
    regions = None if bedfile in (None, "False", False) else regions_from_bed(bedfile, header)
    packed = fastpack.pack_reads(reads, "sscs", bdelim, regions)
    t_pack = time.time()
    own = engine is None
    engines = [engine] if engine is not None else [GpuConsensus(device + k, cutoff=cutoff, scorrect=scorrect)
                                                   for k in range(max(1, int(gpus)))]
    try:
        for eng in engines:
            eng.set_params(cutoff=cutoff, scorrect=scorrect)
        res, packed, n_chunks = stream.consensus_chunked(packed, engines, regions, chunk_reads, log=log)
    finally:
        if own:
            for eng in engines:
                eng.close()
    t_gpu = time.time()
    os.makedirs(os.path.join(sample_dir, "sscs"))
    # record assembly, the sort / merge steps between the stages and the BAM + BAI writing are native
    # host code (csrc/ccemit.cpp); emit.emit_all + bam.write_bam are the same thing in Python
    where = {"sscs.sorted": "sscs", "singleton.sorted": "sscs", "badReads": "sscs",
             "dcs.sorted": "dcs", "sscs.singleton.sorted": "dcs",
             "sscs.correction.sorted": "sscs_sc", "singleton.correction.sorted": "sscs_sc",
             "uncorrected.sorted": "sscs_sc", "sscs.sc.sorted": "sscs_sc",
             "dcs.sc.sorted": "dcs_sc", "sscs.sc.singleton.sorted": "dcs_sc", "all.unique.dcs.sorted": "dcs_sc"}
    sc_only = {"sscs_sc", "dcs_sc"}
    paths = {}
    for name, sub in where.items():
        if sub in sc_only and not scorrect:
            continue
        os.makedirs(os.path.join(sample_dir, sub), exist_ok=True)
        paths[name] = os.path.join(sample_dir, sub, "%s.%s.bam" % (identifier, name))
    _, stats, fam_txt = fastpack.emit_files(reads, packed, res, scorrect, paths)
    with open(os.path.join(sample_dir, identifier + ".stats.txt"), "w") as fh:
        fh.write(stats)
    with open(os.path.join(sample_dir, identifier + ".read_families.txt"), "w") as fh:
        fh.write(fam_txt)
    t_end = time.time()
    with open(os.path.join(sample_dir, identifier + ".time_tracker.txt"), "w") as fh:
        fh.write("decode+pack: %s\nGPU consensus (%d chunks, %d GPUs): %s\nemit+write: %s\n" %
                 ((t_pack - t0) / 60, n_chunks, len(engines), (t_gpu - t_pack) / 60, (t_end - t_gpu) / 60))
    if cleanup:
        _cleanup(sample_dir, identifier, scorrect)
    log(stats)
    return sample_dir


def _cleanup(sample_dir, identifier, scorrect):
    """ConsensusCruncher.py:306-322."""
    def rm(rel):
        p = os.path.join(sample_dir, rel)
        if os.path.exists(p):
            os.remove(p)
    rm("%s.time_tracker.txt" % identifier)
    rm("sscs/%s.badReads.bam" % identifier)
    for rel in ("dcs/%s.sscs.singleton.sorted.bam" % identifier,):
        rm(rel); rm(rel + ".bai")
    if scorrect:
        for nm in ("sscs_sc/%s.singleton.correction.sorted.bam", "sscs_sc/%s.sscs.correction.sorted.bam",
                   "sscs_sc/%s.uncorrected.sorted.bam", "dcs_sc/%s.sscs.sc.singleton.sorted.bam"):
            rm(nm % identifier); rm(nm % identifier + ".bai")


def build_parser(config_path=None):
    main_p = argparse.ArgumentParser(prog="ConsensusCruncher.py",
                                     description="B200 implementation of ConsensusCruncher's consensus mode")
    main_p.add_argument("-c", "--config", default=None, help="config file; command line options override it")
    sub = main_p.add_subparsers(dest="subparser_name")
    sub.add_parser("fastq2bam", help="not part of the GPU path (needs bwa / picard); use the reference")
    p = sub.add_parser("consensus", help="SSCS, DCS and singleton correction of a UMI-tagged BAM")
    defaults = {"scorrect": "True", "genome": "hg19", "bedfile": None, "cutoff": 0.7, "bdelim": "|",
                "cleanup": None, "bam": None, "c_output": None, "samtools": None}
    if config_path:
        cfg = configparser.ConfigParser()
        cfg.read(config_path)
        if cfg.has_section("consensus"):
            defaults.update(dict(cfg.items("consensus")))
    p.add_argument("-i", "--input", metavar="BAM", dest="bam", type=str, help="input BAM with barcodes in the read names")
    p.add_argument("-o", "--output", metavar="OUTPUT", dest="c_output", type=str, help="output directory")
    p.add_argument("-s", "--samtools", metavar="SAMTOOLS", type=str, help="accepted for compatibility; not used")
    p.add_argument("--scorrect", choices=["True", "False"], help="singleton correction, default: True")
    p.add_argument("-g", "--genome", metavar="VERSION", dest="genome", choices=["hg19", "hg38"])
    p.add_argument("-b", "--bedfile", type=str, help="bed file that divides the genome, or 'False'")
    p.add_argument("--cutoff", type=float, help="consensus cut-off, default 0.7")
    p.add_argument("-d", "--bdelim", metavar="DELIMITER", type=str, help="delimiter before the barcode, default '|'")
    p.add_argument("--cleanup", choices=["True", "False"], help="remove intermediate files")
    p.add_argument("--device", type=int, default=0, help="first CUDA device")
    p.add_argument("--gpus", type=int, default=1, help="GPUs that take the region chunks in turn (devices --device ...)")
    p.add_argument("--compression", type=int, default=None, choices=range(0, 10), metavar="0-9",
                   help="zlib level of the output BAMs (default 6, like samtools; lower is faster, same records)")
    p.add_argument("--chunk-reads", type=int, default=20_000_000,
                   help="reads per chunk of consecutive bed regions (bounds the device memory a sample needs)")
    p.set_defaults(**defaults)
    return main_p, p


def main(argv=None):
    argv = list(sys.argv[1:] if argv is None else argv)
    pre = argparse.ArgumentParser(add_help=False)
    pre.add_argument("-c", "--config", default=None)
    known, _ = pre.parse_known_args(argv)
    main_p, cons_p = build_parser(known.config)
    args = main_p.parse_args(argv)
    if args.subparser_name is None:
        main_p.print_help()
        return 0
    if args.subparser_name == "fastq2bam":
        print("fastq2bam is outside the GPU path (bwa + picard); use the reference implementation", file=sys.stderr)
        return 2
    if args.bam is None or args.c_output is None:          # ConsensusCruncher.py:481 (samtools no longer required)
        cons_p.print_help()
        return 0
    if not os.path.isfile(args.bam):
        cons_p.error("input BAM %r does not exist" % args.bam)
    bedfile = args.bedfile
    if args.genome == "hg38":                              # overrides even '-b False' (ConsensusCruncher.py:143-145)
        bedfile = default_bedfile("hg38")
        if bedfile is None:
            cons_p.error("-g hg38 needs hg38_cytoBand.txt (see consensuscruncher_b200/cli.py:default_bedfile)")
    elif bedfile is None:                                  # documented default: the hg19 cytoband (README.md:77-80)
        bedfile = default_bedfile("hg19")
        if bedfile is None:
            cons_p.error("no -b/--bedfile given and hg19_cytoBand.txt is not installed; "
                         "pass -b FILE or -b False")
    from .engine import GpuError
    if args.compression is not None:
        os.environ["CCP_DEFLATE_LEVEL"] = str(args.compression)
    try:
        run_consensus(args.bam, args.c_output, cutoff=float(args.cutoff), scorrect=str(args.scorrect) != "False",
                      bedfile=bedfile, bdelim=args.bdelim, cleanup=str(args.cleanup) == "True", device=args.device,
                      gpus=args.gpus, chunk_reads=args.chunk_reads)
    except (pack.DomainError, GpuError, ValueError, FileExistsError) as exc:
        # one line, no traceback, nothing half-written: the sample directory is only created once the GPU pass is done
        print("ConsensusCruncher.py consensus: %s: %s" % (type(exc).__name__, exc), file=sys.stderr)
        return 1
    return 0


if __name__ == "__main__":
    sys.exit(main())

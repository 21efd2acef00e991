"""Driver with the reference's surface (nanotensor/run_nanotensor.py): ``CommandLine`` (``-c
config.json``), ``RunTensorflow(args).run_model() / test_model()`` and ``main()``, consuming the
UNMODIFIED config JSONs.  Training (``"train": true``) is outside the B200 forward path and
raises ``Usage``.

Multi-GPU: the reference builds one tower per free GPU, each pulling the next batch
(run_nanotensor.py:188-202).  Here reads are sharded across ranks (one process per GPU,
``shard.assign_reads``), no collective on the data path; see bench.py for the torchrun launch.
"""
from __future__ import print_function

import argparse
import logging as log
import os
import sys
from timeit import default_timer as timer

import numpy as np

from .checkpoint import latest_checkpoint
from .dataset import FullSignalSequence, MotifSequence, NumpyEventData, OutOfRangeError
from .error import Usage
from .network import CrossEntropy, CtcLoss
from .shard import assign_reads
from .utils import DotDict, list_dir, load_json, merge_two_dicts, time_it


class CommandLine(object):
    """run_nanotensor.py:36-120."""

    def __init__(self, in_args=None):
        self.parser = argparse.ArgumentParser(
            description='Run the NanoTensor BLSTM basecaller forward on B200 from a JSON config',
            usage='%(prog)s -c path/to/config.json')
        self.parser.add_argument('-c', '--config', help='path to a json config file', required=True)
        self.parser.add_argument('-v', '--verbose', action='store_true', help='verbose option')
        self.args = vars(self.parser.parse_args(in_args))

    @staticmethod
    def check_args(args):
        """Validate a config dict as run_nanotensor.py:94-120 does; returns a DotDict."""
        args = DotDict(args)
        modes = [bool(args.get("train")), bool(args.get("inference")), bool(args.get("test"))]
        if sum(modes) != 1:
            raise Usage("Need to specify exactly one of train, inference or test in the config")
        if args.get("train"):
            raise Usage("training is outside the B200 forward path: set inference or test")
        cd = args.get("CreateDataset")
        if not cd or "dataset" not in cd:
            raise Usage("config needs CreateDataset.dataset")
        if args.get("inference"):
            d = cd.get("inference_dir")
            if not d or not os.path.isdir(d):
                raise Usage("inference_dir does not exist: {}".format(d))
            out = cd.get("dataset_args", {}).get("inference_output_dir")
            if out and not os.path.isdir(out):
                raise Usage("inference_output_dir does not exist: {}".format(out))
        if args.get("test"):
            d = cd.get("test_dir")
            if not d or not os.path.isdir(d):
                raise Usage("test_dir does not exist: {}".format(d))
        if not args.get("network"):
            raise Usage("config needs a network list")
        if args.get("use_checkpoint"):
            if not args.get("trained_model") or not os.path.isdir(args["trained_model"]):
                raise Usage("trained_model directory does not exist: {}".format(args.get("trained_model")))
        elif not args.get("trained_model_path") or not os.path.exists(args["trained_model_path"] + ".index"):
            raise Usage("trained_model_path does not exist: {}".format(args.get("trained_model_path")))
        return args


class RunTensorflow(object):
    """run_nanotensor.py:123-396, inference and test routes."""

    dataset_options = {"FullSignalSequence": FullSignalSequence, "MotifSequence": MotifSequence,
                       "NumpyEventData": NumpyEventData}
    graph_options = {"CtcLoss": CtcLoss, "CrossEntropy": CrossEntropy}

    def __init__(self, args, rank=0, world_size=1, device=None):
        self.args = args if isinstance(args, DotDict) else DotDict(args)
        try:
            self.Dataset = self.dataset_options[self.args.CreateDataset.dataset]
        except KeyError:
            raise Usage("unsupported dataset {!r}".format(self.args.CreateDataset.dataset))
        try:
            self.Graph = self.graph_options[self.args.BuildGraph.graph]
        except KeyError:
            raise Usage("unsupported graph {!r}".format(self.args.BuildGraph.graph))
        self.rank, self.world_size = rank, world_size
        self.device = device or "cuda:%d" % int(os.environ.get("LOCAL_RANK", 0))
        if self.args.get("train"):
            raise Usage("training is outside the B200 forward path")
        elif self.args.get("test"):
            self.test_files = list_dir(self.args.CreateDataset.test_dir)
            self.testing_opts = []
        elif self.args.get("inference"):
            self.inference_files = list_dir(self.args.CreateDataset.inference_dir, ext=self.args.get("file_ext", ""))
            self.inference_opts = []
        else:
            raise Usage("Need to specify exactly one of train, inference or test in the config")
        log.info("Data Loading Started")
        speed = time_it(self.load_data)
        log.info("Data Loading took {} seconds to complete".format(speed))
        if self.args.get("use_checkpoint"):
            self.model_path = latest_checkpoint(self.args.trained_model)
        else:
            self.model_path = self.args.trained_model_path
        log.info("Initializing Model")
        speed = time_it(self.initialize_model)
        log.info("Model Initialization took {} seconds to complete".format(speed))

    def load_data(self):
        dargs = dict(self.args.CreateDataset.get("dataset_args", {}))
        if self.args.get("test"):
            self.testing = self.Dataset(self.test_files, **merge_two_dicts(
                dargs, {"mode": 1, "verbose": bool(self.args.get("verbose", False))}))
        if self.args.get("inference"):
            self.inference = self.Dataset(self.inference_files, **merge_two_dicts(
                dargs, {"mode": 2, "verbose": bool(self.args.get("verbose", False))}))

    def initialize_model(self):
        self.create_model()

    def create_model(self):
        kw = dict(network=self.args.network, model_path=self.model_path, device=self.device,
                  model_name=self.args.get("model_name"))
        if self.args.get("inference"):
            self.inference_model = self.Graph(dataset=self.inference, summary_name="Inference", **kw)
            self.inference_opts.append(self.inference_model.prediction)
            return "Inference"
        self.testing_model = self.Graph(dataset=self.testing, summary_name="Testing", **kw)
        self.testing_opts.append(self.testing_model.accuracy)
        self.testing_opts.append(self.testing_model.batch_size)
        return "Testing"

    # ------------------------------------------------------------------ inference
    def run_model(self, per_batch=False):
        """run_nanotensor.py:330-353.  ``per_batch=True`` walks the iterator one sess.run-sized batch
        at a time exactly as the reference loop does; the default processes each rank's share of
        the reads as one BN-group-annotated job (identical outputs, large launches)."""
        files = self.inference_files
        if self.world_size > 1:
            counts = [max(1, os.path.getsize(f)) for f in files]
            mine = assign_reads(counts, self.world_size)[self.rank]
            files = [files[i] for i in mine]
        outputs = []
        if per_batch:
            for file_path in files:
                self.inference.file_path = file_path
                self.inference.iterator.initializer()
                prediction_list = []
                try:
                    while True:
                        evaluate_pred = self.inference_model.run([self.inference_opts])
                        prediction_list.extend(evaluate_pred)
                except OutOfRangeError:
                    outputs.append(self.inference.process_output(prediction_list, file_path))
                    log.info(file_path)
        elif type(self.inference) is FullSignalSequence and self.inference_model.graph_name == "CtcLoss":
            # the device-resident chain (job.ReadJob): file bytes -> ntb_parse_signal -> ntb_pack_windows ->
            # ntb_model_forward -> ntb_assemble_reads; the text crosses PCIe once, the called reads come back
            # (same FASTA as dataset.process_output, dataset.py:421-432)
            from .job import ReadJob
            job = ReadJob(self.inference_model.model, step=self.inference.step, start_index=self.inference.start_index,
                          batch_size=self.inference.batch_size)
            for file_path, seq in zip(files, job.run_signal_files(files) if files else []):
                name = os.path.splitext(os.path.basename(file_path))[0]
                path = os.path.join(self.inference.inference_output_dir, name + ".fasta")
                with open(path, "w+") as fasta_f:
                    fasta_f.write(">{}\n{}\n".format(name, seq))
                outputs.append(path)
                log.info(file_path)
        else:
            x, lens, goff, read_off = self.inference.inference_job(files)
            preds, _ = self.inference_model.run_job(x, lens, goff)
            group_read = np.searchsorted(read_off, goff[:-1], side="right") - 1
            for r, file_path in enumerate(files):
                mine = [preds[g] for g in np.nonzero(group_read == r)[0]]
                outputs.append(self.inference.process_output(mine, file_path))
                log.info(file_path)
        print("Finished Inference", file=sys.stderr)
        return outputs

    # ------------------------------------------------------------------ test
    def test_model(self):
        """run_nanotensor.py:355-389: batch-size weighted mean of the per-batch accuracy."""
        x, lens, goff, y = self.testing.test_job()
        preds, _ = self.testing_model.run_job(x, lens, goff)
        acc_sum, step = 0.0, 0
        T = self.testing.seq_len
        for g, pred in enumerate(preds):
            a, b = int(goff[g]), int(goff[g + 1])
            acc = float(np.mean(pred == y[a:b].reshape(-1)))
            acc_sum += acc * (b - a)
            step += (b - a)
            log.info("Iter " + str(step) + ", Testing Accuracy= " + "{:.5f}".format(acc))
        print("End of dataset", file=sys.stderr)
        final = acc_sum / step * 100 if step else float("nan")
        print("Average Accuracy = {:.3f}".format(final), file=sys.stderr)
        return final


def main(in_args=None):
    """run_nanotensor.py:530-576."""
    start = timer()
    command_line = CommandLine(in_args)
    try:
        config = load_json(command_line.args["config"])
        args = CommandLine.check_args(config)
        if args.get("verbose") or command_line.args["verbose"]:
            log.basicConfig(format='%(levelname)s:%(message)s', level=log.DEBUG)
        rank = int(os.environ.get("RANK", 0))
        world = int(os.environ.get("WORLD_SIZE", 1))
        runner = RunTensorflow(args, rank=rank, world_size=world)
        result = runner.run_model() if args.get("inference") else runner.test_model()
        print("Running time = {} seconds".format(timer() - start), file=sys.stderr)
        return result
    except Usage as err:
        command_line.parser.print_usage(sys.stderr)
        print("{}".format(err.msg), file=sys.stderr)
        return 2


if __name__ == "__main__":
    _rc = main()                      # ONCE: the job must not run a second time for the exit code
    sys.exit(_rc if isinstance(_rc, int) else 0)

"""benchmark_bin of this framework: the flags of the reference's lite/api/benchmark.cc:38-73 (same names, same
defaults, same result line), run through the Predictor API of predictor.py on one B200.

    python paddle-lite_b200/benchmark.py --model_dir /path/to/humanseg_shfv75_0119_quant \
        --input_shape 32,3,192,192 --warmup 10 --repeats 100 --result_filename result.txt [--profile]

--optimized_model_path takes the prefix written by PaddlePredictor.SaveOptimizedModel.  --threads and
--power_mode are accepted and ignored (ARM host settings).  --profile prints the profiler's concise and
detailed summaries (lite/core/profile/profiler.cc) from per-instruction CUDA events."""
import argparse
import importlib.util
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))


def _load_package():
    name = "paddle_lite_b200"
    if name in sys.modules:
        return sys.modules[name]
    spec = importlib.util.spec_from_file_location(name, os.path.join(HERE, "__init__.py"), submodule_search_locations=[HERE])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--optimized_model_path", default="", help="the path of the model that is optimized by opt")
    ap.add_argument("--model_dir", default="", help="the path of the model, the model and param files is under model_dir")
    ap.add_argument("--model_filename", default="")
    ap.add_argument("--param_filename", default="")
    ap.add_argument("--input_shape", default="1,3,224,224", help="such as 1,3,244,244")
    ap.add_argument("--input_img_path", default="", help="if not set, the input of model will be 1.0")
    ap.add_argument("--warmup", type=int, default=0)
    ap.add_argument("--repeats", type=int, default=1)
    ap.add_argument("--power_mode", type=int, default=3)
    ap.add_argument("--threads", type=int, default=1)
    ap.add_argument("--result_filename", default="result.txt")
    ap.add_argument("--show_output", action="store_true")
    ap.add_argument("--profile", action="store_true", help="per-instruction timers + profiler summary (eager run)")
    ap.add_argument("--no_cuda_graph", action="store_true", help="launch instruction by instruction")
    ap.add_argument("--device_id", type=int, default=0)
    a = ap.parse_args(argv)
    path = a.optimized_model_path or a.model_dir
    if not path:
        print("Input error, the model path should not be empty.\n")
        ap.print_help()
        return 0
    _load_package()
    from paddle_lite_b200 import predictor
    shape = tuple(int(v) for v in a.input_shape.split(","))
    rec = predictor.benchmark(path, shape, a.warmup, a.repeats, a.result_filename, a.show_output, a.input_img_path,
                              profile=a.profile, cuda_graph=not a.no_cuda_graph, device_id=a.device_id)
    for key in ("summary", "summary_detailed"):
        if key in rec:
            print(rec.pop(key))
    print(json.dumps(rec))
    return 0


if __name__ == "__main__":
    sys.exit(main())

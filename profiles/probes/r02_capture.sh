#!/bin/bash
# Round-2 evidence pass (run on the GPU box through gpurun): the bench line, the launch list of the same command, per-workload
# launch lists, and one `--set full` capture per dominant kernel. Everything lands in gpurun_out/r02/.
O=gpurun_out/r02
mkdir -p $O
python bench.py --steps 20 --warmup 5 > $O/bench.json 2> $O/bench.err || { tail -5 $O/bench.err; exit 1; }
NCU="ncu --clock-control none"
$NCU --metrics gpu__time_duration.sum -c 800 --csv --log-file $O/launches_bench.csv \
    python bench.py --steps 4 --warmup 3 --min-seconds 0 --no-configs --no-cpu-baseline > $O/ncu_bench.log 2>&1
for w in detectron pointpillars centerpoint; do   # our kernels only: the workload set-up launches hundreds of torch kernels first
    $NCU --metrics gpu__time_duration.sum -k "regex:preprocess|det_|pp_|rot_|cp_|vox_|pillar_" -c 120 --csv --log-file $O/launches_$w.csv python profiles/probes/prof_posts.py $w > $O/ncu_$w.log 2>&1
done
$NCU --metrics gpu__time_duration.sum -k regex:jpeg -c 12 --csv --log-file $O/launches_jpeg.csv python profiles/probes/prof_jpeg.py 2 > $O/ncu_jpeg.log 2>&1
full() {  # name, kernel regex, launch-skip, target...
    local name=$1 k=$2 s=$3; shift 3
    $NCU --set full --import-source on -k "regex:$k" -s $s -c 1 -f -o $O/full_$name "$@" > $O/full_$name.log 2>&1
}
full pre        'preprocess'           6  python bench.py --steps 4 --warmup 3 --min-seconds 0 --no-configs --no-cpu-baseline
full filter     'yolo_filter'      6  python bench.py --steps 4 --warmup 3 --min-seconds 0 --no-configs --no-cpu-baseline
full nms        'yolo_nms_kernel'  6  python bench.py --steps 4 --warmup 3 --min-seconds 0 --no-configs --no-cpu-baseline
full detcand    'det_candidates'       2  python profiles/probes/prof_posts.py detectron
full detlevel   'det_level'            2  python profiles/probes/prof_posts.py detectron
full ppcand     'pp_candidates'        2  python profiles/probes/prof_posts.py pointpillars
full decorate   'pillar_decorate'      2  python profiles/probes/prof_posts.py pointpillars
full voxcell    'vox_cell'             2  python profiles/probes/prof_posts.py pointpillars
full voxplace   'vox_place'            2  python profiles/probes/prof_posts.py pointpillars
full voxscatter 'vox_scatter'          2  python profiles/probes/prof_posts.py pointpillars
full rotmask    'rot_mask'             2  python profiles/probes/prof_posts.py pointpillars
full rotpairs   'rot_pairs'            2  python profiles/probes/prof_posts.py pointpillars
full dethist    'det_hist'             2  python profiles/probes/prof_posts.py detectron
full detpost    'det_post'             2  python profiles/probes/prof_posts.py detectron
full flat       'preprocess_flat'      2  python profiles/probes/prof_posts.py detectron
full huff       'jpeg_huffman'         1  python profiles/probes/prof_jpeg.py 1
full cphead     'cp_head'              2  python profiles/probes/prof_posts.py centerpoint
ls -la $O

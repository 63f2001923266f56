# last pass of round 2: round-end style check (suite, smoke, bench, reference arm), then the CLI wall clock with 2 MiB raw pieces
bash scripts/run_final.sh r02z
timeout 600 python scripts/cli_e2e.py --genomes 1000 --repeat 2 > gpurun_out/r02z_cli_e2e_1000.json 2> gpurun_out/r02z_cli.err; tail -c 1500 gpurun_out/r02z_cli_e2e_1000.json; tail -n 3 gpurun_out/r02z_cli.err

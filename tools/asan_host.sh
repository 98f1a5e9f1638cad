#!/bin/bash
# tools/asan_host.sh INDEX_DIR READS: the host side (index loader, FASTQ reader, batcher with its splice / streaming copies /
# line records, SAM writer, meta_analysis, API, CLI) and the kernel source compiled for the host, all under AddressSanitizer +
# UBSan (test infrastructure; no GPU needed).  Builds into gpurun_out/asan/, then runs the CLI in three formats under several
# batching knobs and the API driver tests/hostsim/api_drive.c (read_classify + meta_analysis from the hit records and from a
# copy of the text, three calls on one thread_id).  Prints md5s and the number of sanitizer reports (must be 0).
set -e
cd "$(dirname "$0")/../desamba-so_b200"
O=../gpurun_out/asan; mkdir -p $O
SAN="-fsanitize=address,undefined -fno-omit-frame-pointer -g -O1"
for f in dsb_index dsb_fastq dsb_history dsb_sam dsb_batch dsb_meta dsb_api dsb_main; do
  gcc -std=gnu99 $SAN -I ../include -I host -c -o $O/$f.o host/$f.c
done
g++ -std=c++17 $SAN -Wno-unknown-pragmas -fopenmp -I ../include -I csrc -c -o $O/hostsim.o ../tests/hostsim/dsb_hostsim.cpp
g++ -std=c++17 $SAN -Wno-unknown-pragmas -fopenmp -I ../include -I csrc -c -o $O/buildsim.o ../tests/hostsim/dsb_build_hostsim.cpp
gcc -std=gnu99 $SAN -I ../include -c -o $O/api_drive.o ../tests/hostsim/api_drive.c
OBJ="$O/dsb_index.o $O/dsb_fastq.o $O/dsb_history.o $O/dsb_sam.o $O/dsb_batch.o $O/dsb_meta.o $O/dsb_api.o $O/hostsim.o $O/buildsim.o"
g++ -fsanitize=address,undefined -fopenmp -o $O/api_asan $O/api_drive.o $OBJ -lz -lm -lpthread
g++ -fsanitize=address,undefined -fopenmp -o $O/cli_asan $O/dsb_main.o $OBJ -lz -lm -lpthread
IDX=$1; R=$2
for env in "X=1" "DSB_BATCH_READS=37" "DSB_SCAN_PIECE=1048576" "DSB_NT_COPY=0" "DSB_FORMAT_THREADS=1"; do
  for fmt in SAM SAM_FULL DES; do
    env $env $O/cli_asan classify -f $fmt $IDX/ $R 2> $O/asan.err | md5sum | cut -c1-12 | tr '\n' ' '
    grep -c "ERROR\|runtime error" $O/asan.err | tr '\n' ' '
  done
  echo " <- $env (md5 + sanitizer reports per format: SAM, SAM_FULL, DES)"
done
DSB_SCAN_PIECE=1048576 DSB_BATCH_BASES=700000 DSB_SLOTS=3 $O/api_asan $IDX/ $R 2>&1 | tail -4

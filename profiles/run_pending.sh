# applies profiles/pending_loop_block_kernel_arg.patch ON THE BOX (the tree at home stays at the tested state), rebuilds, and runs the
# quick check: GPU parity tests, bench line, loop trace, mapping loop
set -e
patch -p1 < profiles/pending_loop_block_kernel_arg.patch
python -c "import __graft_entry__ as g; g.build()" > gpurun_out/build_pending.log 2>&1 || { tail -20 gpurun_out/build_pending.log; exit 1; }
set +e
bash profiles/run_quick.sh

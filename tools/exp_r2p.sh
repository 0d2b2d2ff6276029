# the allocation-free entry point: its test alone in a fresh process, then the whole GPU suite (the file collects last)
python -c "import __graft_entry__ as g; g.build()"
echo "== alone"; python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -4
echo "== after the other tests"; python -m pytest tests -m gpu -x -q 2>&1 | tail -4
echo "== compute-sanitizer memcheck on the new test"; timeout 600 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -5

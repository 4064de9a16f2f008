set -u
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python -c "import __graft_entry__ as g; g.smoke()"
VO_PROBE_PARITY=1 timeout 300 python tools/noise_probe.py 0 2 8 2>&1 | grep -E "noise|parity|Error" | sed 's/features.*per 64 pairs://'

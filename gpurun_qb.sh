python -m pytest tests -x -q -m gpu 2>&1 | grep -E "^E|passed|failed|Error" | head -14 | cut -c1-300

"""Incremental-state keys in the reference's scheme "{ClassName}.{instance id}.{key}" (ttl/tell/utils/state.py:10-34:
one counter per class name, an id handed to every module instance on first use), so that callers written against the
reference -- `decoder.filter_incremental_state(state, active_idx)`, `reorder_incremental_state(state, new_order)`,
code that looks for 'DynamicConv1dTBC' in the keys (decoder_flattened.py:202-207) -- keep working on the state the
B200 mirror stores in the same dict."""
from collections import defaultdict

_NEXT_ID = defaultdict(int)


def new_instance_id(class_name):
    _NEXT_ID[class_name] += 1
    return _NEXT_ID[class_name]


def full_key(class_name, instance_id, key):
    return '%s.%d.%s' % (class_name, instance_id, key)

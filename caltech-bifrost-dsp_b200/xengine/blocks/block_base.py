"""Block base class: runtime-control key registry, validation, pending -> active swap, stats.

Mirror of the public surface of the reference's ``Block`` (pipeline/lwa352_pipeline/blocks/block_base.py:22-391):
same constructor arguments, same method names and return codes, same JSON command envelope
``{"id": ..., "cmd": "update", "val": {"kwargs": {...}}}`` (block_base.py:216-255) and the same etcd key naming
(block_base.py:127-144), so scripts and control clients written against the reference keep working.  etcd
itself is out of scope: any object with ``add_watch_prefix_callback`` / ``put`` / ``cancel_watch`` is accepted.
"""
import json
import socket
import threading
import time

from .. import runtime as rt

COMMAND_OK = 0
COMMAND_NOT_RECOGNIZED = -1
COMMAND_WRONG_TYPE = -2
COMMAND_INVALID = -3

_PROCLOGS = ("bind", "in", "out", "size", "sequence0", "perf", "stats")


class _Event(object):
    def __init__(self, value):
        self.value = value


class _WatchResponse(object):
    def __init__(self, values):
        self.events = [_Event(v) for v in values]


class Block(object):
    pipeline_id = 0
    _instance_count = -1

    @classmethod
    def set_id(cls, x):
        cls.pipeline_id = x

    @classmethod
    def _get_instance_id(cls):
        # one counter per subclass (the attribute is created on the subclass by the augmented assignment)
        cls._instance_count += 1
        return cls._instance_count

    def __init__(self, log, iring, oring, guarantee, core, etcd_client=None, command_keyroot='/cmd/corr',
                 monitor_keyroot='/mon/corr', response_keyroot='/resp/corr', name=None):
        self.log, self.iring, self.oring = log, iring, oring
        self.guarantee, self.core = guarantee, core
        self.instance_id = self._get_instance_id()
        self.name = name or type(self).__name__
        self.stats = {}
        self.log.info("Pipeline %d: Initializing block: %s (instance %d)"
                      % (self.pipeline_id, self.name, self.instance_id))

        cls_name = type(self).__name__
        for short in _PROCLOGS:
            attr = "sequence_proclog" if short == "sequence0" else short + "_proclog"
            setattr(self, attr, rt.ProcLog("%s/%s" % (cls_name, short)))
        self.in_proclog.update({'nring': 1, 'ring0': getattr(self.iring, 'name', None)})
        if self.oring is not None:
            self.out_proclog.update({'nring': 1, 'ring0': getattr(self.oring, 'name', None)})

        self.etcd_client = etcd_client
        suffix = "/x/%s/pipeline/%s/%s/%s" % (socket.gethostname(), self.pipeline_id, self.name, self.instance_id)
        self.command_key = command_keyroot + suffix
        self.monitor_key = monitor_keyroot + suffix
        self.response_key = response_keyroot + suffix

        self._control_lock = threading.Lock()
        self._etcd_watch_id = None
        if self.etcd_client:
            self.log.info("Adding watch callback to %s" % self.command_key)
            self._etcd_watch_id = self.etcd_client.add_watch_prefix_callback(self.command_key, self._etcd_callback)

        self.update_pending = False
        self.command_vals = {}
        self._pending_command_vals = {}
        self._command_types = {}
        self._command_conditions = {}
        self._etcd_sets_pending = True

    # ------------------------------------------------------------------ command keys
    def define_command_key(self, name, type=None, condition=None, initial_val=None):
        """Register (or redefine) a runtime-control key; ``type`` is checked with isinstance and ``condition``
        is a predicate on the value.  A falsy ``initial_val`` skips both checks (block_base.py:180-185)."""
        if initial_val:
            if type:
                assert isinstance(initial_val, type), "%s: key %s: Initial value type check fail!" % (self.name, name)
            if condition:
                assert condition(initial_val), "%s: key %s: Intial value failed condition check!" % (self.name, name)
        self.command_vals[name] = initial_val
        self._pending_command_vals[name] = initial_val
        self._command_types[name] = type
        self._command_conditions[name] = condition

    def process_command_strings(self, cmds):
        """Inject JSON command strings as if etcd had delivered them (block_base.py:194-214)."""
        if not isinstance(cmds, list):
            cmds = [cmds]
        self._etcd_callback(_WatchResponse(cmds))

    def _decode_command(self, raw):
        """Envelope checks shared by every block.  Returns (seq_id, kwargs) or (seq_id, error string)."""
        v = json.loads(raw)
        seq_id = v.get('id', None)
        if seq_id is None:
            return "0", "Missing ID field"
        if v.get('cmd', None) != "update":
            return "0", "Invalid command"
        val = v.get("val", None)
        if not isinstance(val, dict):
            return seq_id, "`val` field should be a dictionary"
        kwargs = val.get("kwargs", None)
        if not isinstance(kwargs, dict):
            return seq_id, "`val[kwargs]` field should be a dictionary"
        return seq_id, kwargs

    def _etcd_callback(self, watchresponse):
        rt.set_core(self.core)
        with self._control_lock:
            for event in watchresponse.events:
                seq_id, payload = self._decode_command(event.value)
                if not isinstance(payload, dict):
                    self._send_command_response(seq_id, False, payload)
                    continue
                try:
                    rc = self._process_commands(payload, set_pending_flag=self._etcd_sets_pending)
                except Exception:
                    rc = COMMAND_INVALID
                self.update_stats({'last_cmd_response': rc})
                self._send_command_response(seq_id, rc == COMMAND_OK, str(rc))

    def _send_command_response(self, seq_id, processed_ok, response):
        resp = {'id': seq_id,
                'val': {'status': 'normal' if processed_ok else 'error', 'response': response,
                        'timestamp': time.time()}}
        if self.etcd_client:
            try:
                self.etcd_client.put(self.response_key, json.dumps(resp))
            except Exception:
                self.log.error("Error trying to send ETCD command response")
                raise
        else:
            self.log.info("No ETCD interface: Command response: %s" % (resp,))
        return resp

    def _process_commands(self, command_dict, set_pending_flag=True):
        """Validate and stage ``command_dict`` into the pending set.  Returns COMMAND_OK, or the first failure
        (COMMAND_NOT_RECOGNIZED / COMMAND_WRONG_TYPE / COMMAND_INVALID), in which case ``update_pending`` is
        not raised (keys staged before the failing one stay staged, as in block_base.py:296-347)."""
        for key, val in command_dict.items():
            if key not in self.command_vals:
                self.log.error("%s: Command key %s not recognized" % (self.name, key))
                return COMMAND_NOT_RECOGNIZED
            want = self._command_types[key]
            if want and not isinstance(val, want):
                self.log.error("%s: Command key %s had wrong type (had %s, expected %s)"
                               % (self.name, key, type(val), want))
                return COMMAND_WRONG_TYPE
            cond = self._command_conditions[key]
            if cond and not cond(val):
                self.log.error("%s: Command key %s failed requirements" % (self.name, key))
                return COMMAND_INVALID
            self._pending_command_vals[key] = val
            self.stats['new_' + key] = val
        if set_pending_flag:
            self.update_pending = True
        self.stats['update_pending'] = True
        self.stats['last_cmd_time'] = time.time()
        return COMMAND_OK

    def update_command_vals(self):
        """Promote pending command values to the active set (called from ``main()``)."""
        with self._control_lock:
            self.command_vals.update(self._pending_command_vals)
            self.update_pending = False
            self.stats['update_pending'] = False
            self.stats['last_cmd_proc_time'] = time.time()
        self.update_stats(self.command_vals)

    def acquire_control_lock(self):
        self._control_lock.acquire()

    def release_control_lock(self):
        self._control_lock.release()

    def update_stats(self, new_stats=None):
        """Merge into ``self.stats`` and publish the whole dict (ProcLog.update replaces its contents)."""
        if new_stats:
            self.stats.update(new_stats)
        self.stats_proclog.update(self.stats)

    def __del__(self):
        if getattr(self, "_etcd_watch_id", None):
            self.etcd_client.cancel_watch(self._etcd_watch_id)

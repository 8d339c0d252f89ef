"""OpenAI-wire message normalisation and Hermes tool-call parsing (boundary B).

The reference starts vLLM with `--enable-auto-tool-choice --tool-call-parser hermes`
(`docker-compose.vllm.yml:50-51`): tools are described to the model in a `<tools>` block of the
system prompt, the model answers `<tool_call>{"name": …, "arguments": {…}}</tool_call>`, and the
server turns that text into `delta.tool_calls` entries, which is the shape the reference
consumes (`app/core/vllm_handler.py:394-408`) and sends back as history
(`vllm_handler.py:425-440`: an `assistant` message with `tool_calls`, then a `tool` message).
Nothing here comes from the reference's sources — it has none for this; the format is the
public Hermes function-calling convention.
"""
from __future__ import annotations

import json
import uuid
from typing import Any, Dict, List, Optional, Sequence, Tuple

TOOL_OPEN, TOOL_CLOSE = "<tool_call>", "</tool_call>"

_TOOLS_PREAMBLE = (
    "You are a function calling AI model. You are provided with function signatures within "
    "<tools></tools> XML tags. You may call one or more functions to assist with the user query. "
    "Don't make assumptions about what values to plug into functions. Here are the available tools: "
    "<tools> {tools} </tools>\n"
    "For each function call return a json object with function name and arguments within "
    "<tool_call></tool_call> XML tags as follows:\n"
    "<tool_call>\n{{\"name\": <function-name>, \"arguments\": <args-dict>}}\n</tool_call>"
)


class ChatFormatError(ValueError):
    """A request the chat endpoint must answer with HTTP 400."""

    def __init__(self, message: str, param: Optional[str] = None):
        super().__init__(message)
        self.param = param


def _text_of(content: Any, where: str) -> str:
    """OpenAI `content` is a string, null, or a list of parts; only text parts are served."""
    if content is None:
        return ""
    if isinstance(content, str):
        return content
    if isinstance(content, list):
        out = []
        for part in content:
            if isinstance(part, str):
                out.append(part)
            elif isinstance(part, dict) and part.get("type") in ("text", "input_text", None) and "text" in part:
                out.append(str(part["text"]))
            else:
                raise ChatFormatError(f"{where}: only text content parts are supported", "messages")
        return "".join(out)
    raise ChatFormatError(f"{where}: content must be a string, null or a list of parts", "messages")


def render_tool_call(name: str, arguments: Any) -> str:
    if isinstance(arguments, str):
        try:
            arguments = json.loads(arguments) if arguments.strip() else {}
        except json.JSONDecodeError:
            pass                                      # keep the raw string: the model sees what it wrote
    return f"{TOOL_OPEN}\n{json.dumps({'name': name, 'arguments': arguments}, ensure_ascii=False)}\n{TOOL_CLOSE}"


def normalise_messages(messages: Sequence[Dict[str, Any]], tools: Optional[Sequence[Dict[str, Any]]] = None,
                       tool_choice: Any = None) -> List[Dict[str, str]]:
    """OpenAI chat messages (+ tools) → plain `[{"role", "content"}]` for the Llama-3 template.

    * `tools` go into the system message (created when the conversation has none);
      `tool_choice == "none"` leaves them out.
    * an `assistant` message with `tool_calls` becomes its `<tool_call>` text — byte-identical to
      what the parser below accepted when the model produced it, so the re-sent history still
      matches the session's cached KV prefix.
    * a `tool` message becomes a `tool`-role turn holding `<tool_response>`.
    """
    if not isinstance(messages, (list, tuple)) or not messages:
        raise ChatFormatError("messages must be a non-empty list", "messages")
    out: List[Dict[str, str]] = []
    for i, m in enumerate(messages):
        if not isinstance(m, dict) or "role" not in m:
            raise ChatFormatError(f"messages[{i}] must be an object with a role", "messages")
        role = str(m["role"])
        text = _text_of(m.get("content"), f"messages[{i}]")
        if role == "developer":
            role = "system"
        if role == "assistant" and m.get("tool_calls"):
            calls = []
            for tc in m["tool_calls"]:
                fn = (tc or {}).get("function") or {}
                if not fn.get("name"):
                    raise ChatFormatError(f"messages[{i}].tool_calls: function.name is required", "messages")
                calls.append(render_tool_call(str(fn["name"]), fn.get("arguments", {})))
            text = (text + "\n" if text else "") + "\n".join(calls)
        elif role == "tool":
            text = f"<tool_response>\n{text}\n</tool_response>"
        elif role not in ("system", "user", "assistant"):
            raise ChatFormatError(f"messages[{i}]: unknown role {role!r}", "messages")
        out.append({"role": role, "content": text})
    if tools and tool_choice != "none":
        specs = []
        for j, t in enumerate(tools):
            if not isinstance(t, dict) or t.get("type", "function") != "function" or not (t.get("function") or {}).get("name"):
                raise ChatFormatError(f"tools[{j}] must be a function tool with a name", "tools")
            specs.append(json.dumps({"type": "function", "function": t["function"]}, ensure_ascii=False))
        block = _TOOLS_PREAMBLE.format(tools=" ".join(specs))
        if out[0]["role"] == "system":
            out[0] = {"role": "system", "content": (out[0]["content"] + "\n\n" if out[0]["content"] else "") + block}
        else:
            out.insert(0, {"role": "system", "content": block})
    return out


def new_tool_call_id() -> str:
    return "chatcmpl-tool-" + uuid.uuid4().hex


class HermesToolParser:
    """Incremental splitter of generated text into content and tool calls.

    `feed(text)` / `flush()` return events:
      ("content", str)
      ("tool_call", index, call_id, name, arguments_json_str)
    Text that may still turn out to be the start of `<tool_call>` is held back; the body of a
    call is buffered until `</tool_call>` (or end of stream) and then parsed.  A body that is
    not a JSON object with a `name` is given back as content, tags included, so nothing the
    model wrote is lost.
    """

    def __init__(self, enabled: bool = True):
        self.enabled = enabled
        self.buf = ""
        self.in_call = False
        self.n_calls = 0

    def _parse(self, body: str) -> Optional[Tuple[str, str]]:
        try:
            obj = json.loads(body.strip())
        except json.JSONDecodeError:
            return None
        if not isinstance(obj, dict) or not isinstance(obj.get("name"), str):
            return None
        args = obj.get("arguments", obj.get("parameters", {}))
        return obj["name"], args if isinstance(args, str) else json.dumps(args, ensure_ascii=False)

    def _close(self, body: str, closed: bool) -> List[tuple]:
        parsed = self._parse(body)
        if parsed is None:
            return [("content", TOOL_OPEN + body + (TOOL_CLOSE if closed else ""))]
        ev = ("tool_call", self.n_calls, new_tool_call_id(), parsed[0], parsed[1])
        self.n_calls += 1
        return [ev]

    def feed(self, text: str) -> List[tuple]:
        if not self.enabled:
            return [("content", text)] if text else []
        self.buf += text
        events: List[tuple] = []
        while True:
            if self.in_call:
                k = self.buf.find(TOOL_CLOSE)
                if k < 0:
                    return events
                body, self.buf, self.in_call = self.buf[:k], self.buf[k + len(TOOL_CLOSE):], False
                events += self._close(body, True)
                continue
            k = self.buf.find(TOOL_OPEN)
            if k >= 0:
                head = self.buf[:k]
                # whitespace between content and a call (or between two calls) belongs to neither
                if head.strip():
                    events.append(("content", head))
                self.buf, self.in_call = self.buf[k + len(TOOL_OPEN):], True
                continue
            hold = 0
            for n in range(min(len(TOOL_OPEN) - 1, len(self.buf)), 0, -1):
                if self.buf.endswith(TOOL_OPEN[:n]):
                    hold = n
                    break
            out = self.buf[:len(self.buf) - hold]
            if self.n_calls and not out.strip():
                # only whitespace since the last call: keep holding, it may separate two calls
                return events
            if out:
                events.append(("content", out))
            self.buf = self.buf[len(self.buf) - hold:] if hold else ""
            return events

    def flush(self) -> List[tuple]:
        if not self.enabled:
            return []
        buf, self.buf = self.buf, ""
        if self.in_call:
            self.in_call = False
            return self._close(buf, False)
        if buf and not (self.n_calls and not buf.strip()):
            return [("content", buf)]
        return []

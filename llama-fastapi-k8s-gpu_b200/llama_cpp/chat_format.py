"""Chat templating: messages -> prompt string (+ stop strings).

Restates the formatters llama-cpp-python 0.2.77 uses under `create_chat_completion`
(/root/reference/api.py:55-63; upstream `llama_cpp/llama_chat_format.py`, un-vendored, SURVEY.md
2.2 H4): exact-match guess of the GGUF `tokenizer.chat_template` -> built-in "llama-3" /
"chatml" / "mistral-instruct" formatter, else the template itself through Jinja2, else "llama-2".
The built-in formatters silently drop messages whose role they do not know -- that matters for the
reference, which forwards the client's `turn` field as the role (api.py:122-125).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Callable, Dict, List, Optional, Tuple, Union

CHATML_CHAT_TEMPLATE = ("{% for message in messages %}{{'<|im_start|>' + message['role'] + '\n' + message['content'] + "
                        "'<|im_end|>' + '\n'}}{% endfor %}{% if add_generation_prompt %}{{ '<|im_start|>assistant\n' }}{% endif %}")
MISTRAL_INSTRUCT_CHAT_TEMPLATE = ("{{ bos_token }}{% for message in messages %}{% if (message['role'] == 'user') != "
                                  "(loop.index0 % 2 == 0) %}{{ raise_exception('Conversation roles must alternate "
                                  "user/assistant/user/assistant/...') }}{% endif %}{% if message['role'] == 'user' %}"
                                  "{{ '[INST] ' + message['content'] + ' [/INST]' }}{% elif message['role'] == 'assistant' %}"
                                  "{{ message['content'] + eos_token + ' ' }}{% else %}{{ raise_exception('Only user and "
                                  "assistant roles are supported!') }}{% endif %}{% endfor %}")
LLAMA3_INSTRUCT_CHAT_TEMPLATE = (
    "{% set loop_messages = messages %}{% for message in loop_messages %}{% set content = "
    "'<|start_header_id|>' + message['role'] + '<|end_header_id|>\n\n'+ message['content'] | trim + "
    "'<|eot_id|>' %}{% if loop.index0 == 0 %}{% set content = bos_token + content %}{% endif %}"
    "{{ content }}{% endfor %}{{ '<|start_header_id|>assistant<|end_header_id|>\n\n' }}")


@dataclass
class ChatFormatterResponse:
    prompt: str
    stop: Optional[Union[str, List[str]]] = None
    added_special: bool = False      # True: the prompt already carries BOS, do not add another


def _map_roles(messages: List[Dict[str, Any]], role_map: Dict[str, str]) -> List[Tuple[str, Optional[str]]]:
    out = []
    for m in messages:
        role = m.get("role")
        if role in role_map:
            c = m.get("content")
            out.append((role_map[role], c if isinstance(c, str) else None))
    return out


def _system_message(messages) -> str:
    for m in messages:
        if m.get("role") == "system":
            return m.get("content") or ""
    return ""


def format_llama3(messages, **kw) -> ChatFormatterResponse:
    roles = dict(system="<|start_header_id|>system<|end_header_id|>\n\n",
                 user="<|start_header_id|>user<|end_header_id|>\n\n",
                 assistant="<|start_header_id|>assistant<|end_header_id|>\n\n")
    sep = "<|eot_id|>"
    msgs = _map_roles(messages, roles)
    msgs.append((roles["assistant"], None))
    ret = "<|begin_of_text|>"
    for role, content in msgs:
        ret += role + content + sep if content else role
    return ChatFormatterResponse(prompt=ret, stop=sep)


def format_chatml(messages, **kw) -> ChatFormatterResponse:
    roles = dict(user="<|im_start|>user", assistant="<|im_start|>assistant")
    sep = "<|im_end|>"
    sysm = _system_message(messages)
    sysm = f"<|im_start|>system\n{sysm}" if sysm else ""
    msgs = _map_roles(messages, roles)
    msgs.append((roles["assistant"], None))
    ret = "" if sysm == "" else sysm + sep + "\n"
    for role, content in msgs:
        ret += role + "\n" + content + sep + "\n" if content else role + "\n"
    return ChatFormatterResponse(prompt=ret, stop=sep)


def format_mistral_instruct(messages, **kw) -> ChatFormatterResponse:
    bos, eos = "<s>", "</s>"
    prompt = bos
    for m in messages:
        c = m.get("content")
        if m.get("role") == "user" and isinstance(c, str):
            prompt += "[INST] " + c
        elif m.get("role") == "assistant" and c is not None:
            prompt += " [/INST]" + c + eos
    prompt += " [/INST]"
    return ChatFormatterResponse(prompt=prompt, stop=eos)


def format_llama2(messages, **kw) -> ChatFormatterResponse:
    roles = dict(user="<s>[INST]", assistant="[/INST]")
    msgs = _map_roles(messages, roles)
    sysm = _system_message(messages)
    if sysm:
        sysm = f"<s>[INST] <<SYS>>\n{sysm}\n<</SYS>>"
    seps = [" ", "</s>"]
    ret = sysm + seps[0]
    for i, (role, content) in enumerate(msgs):
        if sysm and i == 0:
            ret += (content or "") + seps[i % 2]
        elif content:
            ret += role + content + " " + seps[i % 2]
        else:
            ret += role + " "
    return ChatFormatterResponse(prompt=ret + "[/INST]")


FORMATS: Dict[str, Callable[..., ChatFormatterResponse]] = {
    "llama-3": format_llama3, "chatml": format_chatml, "mistral-instruct": format_mistral_instruct,
    "llama-2": format_llama2,
}


def guess_chat_format_from_gguf_metadata(metadata: Dict[str, Any]) -> Optional[str]:
    t = metadata.get("tokenizer.chat_template")
    if t is None:
        return None
    if t == CHATML_CHAT_TEMPLATE:
        return "chatml"
    if t == MISTRAL_INSTRUCT_CHAT_TEMPLATE:
        return "mistral-instruct"
    if t == LLAMA3_INSTRUCT_CHAT_TEMPLATE:
        return "llama-3"
    return None


class Jinja2ChatFormatter:
    def __init__(self, template: str, eos_token: str, bos_token: str, add_generation_prompt: bool = True):
        import jinja2
        from jinja2.sandbox import ImmutableSandboxedEnvironment
        self.eos_token, self.bos_token = eos_token, bos_token
        self.add_generation_prompt = add_generation_prompt
        env = ImmutableSandboxedEnvironment(loader=jinja2.BaseLoader(), trim_blocks=True, lstrip_blocks=True)
        self._template = env.from_string(template)

    def __call__(self, messages, **kw) -> ChatFormatterResponse:
        def raise_exception(message):
            raise ValueError(message)

        prompt = self._template.render(messages=messages, eos_token=self.eos_token, bos_token=self.bos_token,
                                       raise_exception=raise_exception,
                                       add_generation_prompt=self.add_generation_prompt)
        return ChatFormatterResponse(prompt=prompt, stop=[self.eos_token], added_special=True)

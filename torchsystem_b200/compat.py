"""Where the orchestration layer comes from.

TorchSystem itself (Aggregate, Depends, Service, Producer/Consumer, Compiler, registry) is pure
Python and stays UNCHANGED: when `torchsystem` is importable it is used as is, and every class in
this package plugs into it (tests/test_host_logic.py runs the reference's own handlers and its
test-suite with this package loaded, in the container that holds the reference tree).

The GPU box has no copy of the reference, so for the benchmark and the `-m gpu` tests this module
carries the few lines of behaviour the hot path touches, written from the documented semantics:

* Aggregate.phase / epoch hook      torchsystem/domain/aggregate.py:113-158
* Depends / Provider / inject       torchsystem/depends.py:26-86
* Service.handler / handle          torchsystem/services/service.py:101-152
* Producer / Consumer / event       torchsystem/services/prodcon.py:101-241
* register / gethash                torchsystem/registry/core.py:42-59, accessors.py:45-68

They are a stand-in for running without the reference installed, not a replacement for it.
"""
from __future__ import annotations

import inspect
import re
from contextlib import ExitStack, contextmanager
from dataclasses import dataclass
from hashlib import md5
from json import dumps
from types import GeneratorType, UnionType
from typing import Any, Callable, Union, get_args, get_origin, get_type_hints

import torch

try:  # the real thing, unchanged
    from torchsystem import Aggregate, Depends                                   # type: ignore
    from torchsystem.depends import Provider                                     # type: ignore
    from torchsystem.services import Service, Producer, Consumer, event          # type: ignore
    from torchsystem.registry import gethash, register                           # type: ignore
    HAVE_TORCHSYSTEM = True
except ImportError:
    HAVE_TORCHSYSTEM = False

    class Aggregate(torch.nn.Module):
        """Root of the training aggregate: `phase` drives train()/eval(), `epoch` assignment fires onepoch()."""

        @property
        def id(self) -> Any:
            raise NotImplementedError('The id property must be implemented.')

        @property
        def phase(self) -> str:
            return 'train' if self.training else 'evaluation'

        @phase.setter
        def phase(self, value: str) -> None:
            self.train() if value == 'train' else self.eval()
            self.onphase()

        def onphase(self) -> None: ...

        def onepoch(self) -> None: ...

        def __setattr__(self, name, value):
            fire = name == 'epoch' and hasattr(self, 'epoch')
            super().__setattr__(name, value)
            if fire:
                self.onepoch()

    class Provider:
        def __init__(self):
            self.dependency_overrides: dict[Callable, Callable] = {}

        def override(self, dependency: Callable, override: Callable) -> None:
            self.dependency_overrides[dependency] = override

    class _Dependency:
        def __init__(self, fn: Callable):
            self.callable = fn

    def Depends(fn: Callable) -> Any:
        return _Dependency(fn)

    @contextmanager
    def _generator_dependency(gen):
        try:
            yield next(gen)
        finally:
            next(gen, None)

    def _inject(provider: Provider, fn: Callable) -> Callable:
        sig = inspect.signature(fn)

        def wrapper(*args, **kwargs):
            bound = sig.bind_partial(*args, **kwargs)
            with ExitStack() as stack:
                for name, par in sig.parameters.items():
                    if name in bound.arguments or not isinstance(par.default, _Dependency):
                        continue
                    dep = provider.dependency_overrides.get(par.default.callable, par.default.callable)
                    value = dep()
                    if isinstance(value, GeneratorType):
                        value = stack.enter_context(_generator_dependency(value))
                    bound.arguments[name] = value
                return fn(*bound.args, **bound.kwargs)
        wrapper.__name__ = getattr(fn, '__name__', 'handler')
        wrapper.__wrapped__ = fn
        return wrapper

    def _service_name(name: str) -> str:      # service.py:75  snake_case -> kebab-case
        return name.replace('_', '-')

    def _event_name(name: str) -> str:        # prodcon.py:82  CamelCase -> kebab-case
        return re.sub(r'(?<!^)(?=[A-Z])', '-', name).lower()

    class Service:
        def __init__(self, provider: Provider | None = None):
            self.provider = provider or Provider()
            self.handlers: dict[str, Callable] = {}

        @property
        def dependency_overrides(self) -> dict:
            return self.provider.dependency_overrides

        def handler(self, fn: Callable) -> Callable:
            injected = _inject(self.provider, fn)
            self.handlers[_service_name(fn.__name__)] = injected
            return injected

        def handle(self, action: str, *arguments: Any) -> Any:
            handler = self.handlers.get(action)
            if not handler:
                raise KeyError(f'Handler not found for action: {action}')
            return handler(*arguments)

    class Consumer:
        def __init__(self, provider: Provider | None = None):
            self.provider = provider or Provider()
            self.handlers: dict[str, list[Callable]] = {}

        def handler(self, fn: Callable) -> Callable:
            first = next(iter(inspect.signature(fn).parameters.values()))
            ann = first.annotation
            if isinstance(ann, str):      # stringified by `from __future__ import annotations`
                ann = get_type_hints(fn)[first.name]
            types = get_args(ann) if get_origin(ann) in (Union, UnionType) else (ann,)
            injected = _inject(self.provider, fn)
            for t in types:
                self.handlers.setdefault(_event_name(t.__name__), []).append(injected)
            return injected

        def consume(self, message: Any) -> None:
            for h in self.handlers.get(_event_name(type(message).__name__), []):
                h(message)

    class Producer:
        def __init__(self):
            self.consumers: list[Consumer] = []

        def register(self, *consumers: Consumer) -> None:
            self.consumers.extend(consumers)

        def dispatch(self, message: Any) -> None:
            for consumer in self.consumers:
                consumer.consume(message)

    def event(cls: type) -> type:
        return dataclass(cls)

    def _describe(arg: Any) -> Any:
        if hasattr(arg, '__model__arguments__'):
            name = getattr(arg, '__model__name__', arg.__class__.__name__)
            args = getattr(arg, '__model__arguments__')
            return {'name': name, 'arguments': args} if args else name
        return arg

    def register(cls: type | str, excluded_args: list[int] | None = None, excluded_kwargs: set[str] | None = None):
        """Record constructor arguments on instances as `__model__arguments__` (identity for gethash). `@register` on a
        class, or `@register('name')` to also set `__model__name__`, the name gethash uses instead of the class name
        (torchsystem/registry/accessors.py:150-174, core.py cls_override_init)."""
        if isinstance(cls, str):
            name = cls
            return lambda klass: _register(klass, excluded_args, excluded_kwargs, name)
        return _register(cls, excluded_args, excluded_kwargs, None)

    def _register(cls: type, excluded_args, excluded_kwargs, name):
        excluded_args, excluded_kwargs = excluded_args or [], excluded_kwargs or set()
        init = cls.__init__
        names = list(inspect.signature(cls).parameters)

        def __init__(self, *args, **kwargs):
            init(self, *args, **kwargs)
            captured = {n: _describe(a) for i, (a, n) in enumerate(zip(args, names)) if i not in excluded_args}
            captured |= {k: _describe(v) for k, v in kwargs.items() if k not in excluded_kwargs}
            object.__setattr__(self, '__model__arguments__', captured)
            if name:
                object.__setattr__(self, '__model__name__', name)
        cls.__init__ = __init__
        return cls

    def gethash(obj: object) -> str:
        if hasattr(obj, '__model__hash__'):
            return getattr(obj, '__model__hash__')
        if not hasattr(obj, '__model__arguments__'):
            raise AttributeError(f'The object {obj} was not registered and does not have a hash')
        name = getattr(obj, '__model__name__', obj.__class__.__name__)
        return md5((name + dumps(getattr(obj, '__model__arguments__'))).encode()).hexdigest()
